set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for SYNC in overlap flat overlap flat; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-cpu --no-extra --pdvc-configs yc2_tsp_pdvc_yc2shape,yc2_newmodel_core_bf16 --pdvc-steps 60 --pdvc-warmup 10 --pdvc-sync $SYNC 2> gpurun_out/r2c_$SYNC.err | tail -c 400
done
tail -3 gpurun_out/r2c_overlap.err

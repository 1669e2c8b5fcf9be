# usage: tools/r2_scale.sh N   -- one bench line at N GPUs the way the driver launches it
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
tail -c 700 gpurun_out/r2_bench_n$N.json; tail -3 gpurun_out/r2_bench_n$N.err

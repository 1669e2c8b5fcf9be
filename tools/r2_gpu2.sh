set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2b_pytest.log
timeout 900 python bench.py --steps 20 --no-cpu > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
tail -c 3000 gpurun_out/r2b_bench.json
cat gpurun_out/r2b_pytest.log
tail -5 gpurun_out/r2b_bench.err

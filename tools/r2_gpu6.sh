cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | grep -E "^FAILED|passed|failed" | cut -c1-200 | head -20
timeout 1500 python bench.py --sweep gpurun_out/r2_sweep.json 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 600 gpurun_out/r2_bench_n1.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2>/dev/null; tail -c 300 gpurun_out/r2_bench_ref.json

cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | grep -E "^FAILED|passed|failed" | cut -c1-200 | head -20
timeout 900 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; tail -c 400 gpurun_out/r2_bench_n1.json
python tools/proj_probe.py > /dev/null 2>&1 && ncu --set full --import-source on --clock-control none -k regex:msda_proj --launch-skip 0 -c 18 -f -o gpurun_out/r2_proj python tools/proj_probe.py > gpurun_out/r2p_ncu3.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2p_smoke.log 2>&1 && ncu --set full --import-source on --clock-control none -k regex:msda_ -c 4 -f -o gpurun_out/r2_small python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2p_ncu4.log 2>&1
tail -1 gpurun_out/r2p_smoke.log

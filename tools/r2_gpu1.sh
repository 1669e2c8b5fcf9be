set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/r2a_pytest.log
tools/r2_ab.sh r1 sm100 > gpurun_out/r2a_ab.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:msda_ --launch-skip 6 -c 2 -f -o gpurun_out/r2a python bench.py --no-cpu --no-extra --no-pdvc --steps 1 --warmup 3 > gpurun_out/r2a_ncu.log 2>&1
cat gpurun_out/r2a_pytest.log gpurun_out/r2a_ab.log

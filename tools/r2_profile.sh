# round-2 profiles: (1) launch list of a bench step, (2) ncu --set full of the lean kernels at the headline, (3) of the small
# kernels at the YC2 decoder shape, (4) of the projection GEMMs.  Each capture only after the same command exited 0 without ncu.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
set -x
python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --no-pdvc > gpurun_out/r2p_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extra --no-pdvc > gpurun_out/r2p_ncu1.log 2>&1
ncu --set full --import-source on --clock-control none -k regex:msda_ --launch-skip 6 -c 2 -f -o gpurun_out/r2_lean python bench.py --no-cpu --no-extra --no-pdvc --steps 1 --warmup 3 > gpurun_out/r2p_ncu2.log 2>&1
python tools/proj_bench.py > gpurun_out/r2_proj_bench.txt 2>&1 || exit 1
ncu --set full --import-source on --clock-control none -k regex:msda_proj --launch-skip 0 -c 6 -f -o gpurun_out/r2_proj python tools/proj_probe.py > gpurun_out/r2p_ncu3.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2p_smoke.log 2>&1 || exit 1
ncu --set full --import-source on --clock-control none -k regex:msda_ -c 4 -f -o gpurun_out/r2_small python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2p_ncu4.log 2>&1
tail -2 gpurun_out/r2p_smoke.log; ls -la gpurun_out/*.ncu-rep

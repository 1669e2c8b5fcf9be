# ncu of the round-2 harness kernels (matching, clip + Adam); the probe first runs without ncu
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python tools/harness_kernels_probe.py > gpurun_out/r2h_plain.log 2>&1 || { tail -5 gpurun_out/r2h_plain.log; exit 1; }
ncu --set full --import-source on --clock-control none -k regex:"msda_lsap|msda_adam|msda_sumsq" --launch-skip 6 -c 4 -f -o gpurun_out/r2_harness python tools/harness_kernels_probe.py > gpurun_out/r2h_ncu.log 2>&1
tail -2 gpurun_out/r2h_plain.log; ls -la gpurun_out/r2_harness.ncu-rep

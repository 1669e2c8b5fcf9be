cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | grep -E "^FAILED|AssertionError: |passed|failed|^E  " | cut -c1-220 | head -40

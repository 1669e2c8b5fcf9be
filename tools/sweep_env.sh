#!/bin/bash
# sweep the lean-kernel planning knobs on the headline workload (fwd/bwd ms per launch)
for p in 1 2 4; do for w in 24 32 48 64 96; do
  r=$(MSDA_PASSES=$p MSDA_WIN_KB=$w python bench.py --no-cpu --no-extra --steps 30 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('fwd %.3f ms  bwd %.3f ms' % (r['forward']['ms_per_launch'], r['ms_per_launch']))")
  echo "passes=$p win_kb=$w : $r"
done; done

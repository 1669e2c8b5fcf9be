#!/bin/bash
# sweep the lean-kernel planning knobs on the headline workload (fwd/bwd ms per launch)
run() { python bench.py --no-cpu --no-extra --steps 30 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); r=d['roofline']; print('fwd %.3f ms  bwd %.3f ms' % (r['forward']['ms_per_launch'], r['ms_per_launch']))"; }
for p in 2 3 4; do for w in 24 28 32 40; do echo "passes=$p win_kb=$w : $(MSDA_PASSES=$p MSDA_WIN_KB=$w run)"; done; done
for b in 0 30 34 40 48; do echo "bwin_kb=$b : $(MSDA_BWIN_KB=$b run)"; done

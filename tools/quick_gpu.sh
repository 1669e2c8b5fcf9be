#!/bin/bash
# GPU parity suite + headline fwd/bwd milliseconds (development loop)
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python bench.py --no-cpu --no-extra --steps 30 2>&1 | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('fwd %.3f ms (%.1f%%)  bwd %.3f ms (%.1f%%)  fwd+bwd frac %.3f' % (r['forward']['ms_per_launch'], 100*r['forward']['frac'], r['ms_per_launch'], 100*r['frac'], r['fwd_bwd']['frac']))"

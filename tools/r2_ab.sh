#!/bin/bash
# usage: tools/r2_ab.sh libA libB ...   (names under dense-video-captioning_b200/lib/, e.g. r1 sm100) -- interleaved A/B of the headline
for rep in 1 2; do for L in "$@"; do for DT in f32 bf16; do
MSDA_LIB=$PWD/dense-video-captioning_b200/lib/libmsda_$L.so timeout 300 python bench.py --no-cpu --no-extra --no-pdvc --steps 50 --dtype $DT $AB_EXTRA 2>/tmp/ab_err.log | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$L $DT fwd %.4f ms (%.3f)  bwd %.4f ms (%.3f)  launches/step %.1f' % (r['forward']['ms_per_launch'], r['forward']['frac'], r['ms_per_launch'], r['frac'], d['gpu_launches']/d['steps']))"
done; done; done; tail -3 /tmp/ab_err.log

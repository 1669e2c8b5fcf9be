for L in devA devB devA devB; do MSDA_LIB=$PWD/dense-video-captioning_b200/lib/libmsda_$L.so python bench.py --no-cpu --no-extra --no-pdvc --steps 50 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$L fwd %.4f ms  bwd %.4f ms' % (r['forward']['ms_per_launch'], r['ms_per_launch']))"; done

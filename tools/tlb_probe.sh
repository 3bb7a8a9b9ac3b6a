# pass-1 time per 2^30 cells against the plane size (x-row stride of its stores), nx = nz = 1024
for ny in 32 64 128 256 512 1024 2048; do
  python bench.py --grid 1024,$ny,1024 --steps 10 --warmup 3 --quick --no-e2e --cpu-n 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); ny=$ny
k=d['kernels_ms']; n=[x for x in k if x.startswith('k1')][0]
print('ny %5d  plane %6d KB  %s %.3f ms  -> %.3f ms per 2^30 cells   (%s)' % (ny, ny*4, n, k[n], k[n]*1024/ny, {a: round(b,3) for a,b in k.items()}))"
done

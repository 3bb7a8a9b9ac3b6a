# pass-1 time per 2^30 cells against the plane size (x-row stride of its stores)
#   tools/tlb_probe.sh [nx] [nz]     (defaults 1024 1024)
nx=${1:-1024}; nz=${2:-1024}
for ny in 64 128 256 512 1024 2048; do
  python bench.py --grid $nx,$ny,$nz --steps 10 --warmup 3 --quick --no-e2e --cpu-n 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); nx,ny,nz=$nx,$ny,$nz
k=d['kernels_ms']; n=[x for x in k if x.startswith('k1')][0]
cells=nx*ny*nz
print('grid %dx%dx%d  plane %6d KB  field %5.2f GB  %s %.3f ms  -> %.3f ms per 2^30 cells   (%s)' % (nx,ny,nz, ny*nz*4//1024, cells*4/1e9, n, k[n], k[n]*2**30/cells, {a: round(b,3) for a,b in k.items()}))"
done

"""Per-pass device timing of seed -> delta(x) (also the command profiled under ncu).
usage: prof_step.py [n | nx,ny,nz] [steps]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from randomfield_b200 import _lib, powertools

g = sys.argv[1] if len(sys.argv) > 1 else "1024"
shape = tuple(int(v) for v in g.split(",")) * (3 if "," not in g else 1)
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
eng = _lib.Engine(*shape, dtype=np.float32)
k, sig = powertools.sigma_table(powertools.load_default_power(), shape, 2.0)
eng.set_power(k, sig, 2.0)
eng.set_timing(True)
cells = float(np.prod(shape))
acc = np.zeros(3)
for it in range(steps + 2):
    eng.generate_delta(123 + it, want_stats=True)
    t = np.array(eng.get_timing())
    if it >= 2:
        acc += t
acc /= steps
s1, s2, c = eng.stats()
gbs = [4 * cells / acc[0] / 1e6, 8 * cells / acc[1] / 1e6, 8 * cells / acc[2] / 1e6]
print("%s: k1 %.3f ms (%.0f GB/s)  k2 %.3f ms (%.0f GB/s)  k3 %.3f ms (%.0f GB/s)  sum %.3f ms  %.0f Mcells/s  std %.6f"
      % ("x".join(map(str, shape)), acc[0], gbs[0], acc[1], gbs[1], acc[2], gbs[2], acc.sum(),
         cells / acc.sum() / 1e3, np.sqrt(s2 / c - (s1 / c) ** 2)))

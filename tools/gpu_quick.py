"""Ad-hoc GPU timing probe (not a test): times the three passes at a given size."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from randomfield_b200 import _lib, powertools
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dt = np.float64 if (len(sys.argv) > 2 and sys.argv[2] == 'f64') else np.float32
shape = (n, n, n)
eng = _lib.Engine(*shape, dtype=dt)
k, sig = powertools.sigma_table(powertools.load_default_power(), shape, 2.0)
eng.set_power(k, sig, 2.0)
s = torch.cuda.Stream()
eng.set_stream(s.cuda_stream)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
with torch.cuda.stream(s):
    for it in range(5):
        ev[0].record(s); eng.generate_xpass(123 + it); ev[1].record(s); eng.finish_yz(None, True); ev[2].record(s)
        s.synchronize()
        t1, t2 = ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2])
        cells = n ** 3
        print("iter %d: xpass %.3f ms  yz %.3f ms  total %.3f ms  %.0f Mcells/s  %.0f GB/s(20B/cell)" % (
            it, t1, t2, t1 + t2, cells / (t1 + t2) / 1e3, 20 * cells * (4 if dt == np.float32 else 8) / 4 / (t1 + t2) / 1e6))
s1, s2, c = eng.stats()
print("mean %.3e std %.6f" % (s1 / c, np.sqrt(s2 / c - (s1 / c) ** 2)))

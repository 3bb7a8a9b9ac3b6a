"""Time Generator.measure_power (device r2c + shell binning) at n^3: usage pk_probe.py [n]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import randomfield_b200 as rf
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
g = rf.Generator(n, n, n, 2.0)
g.generate_delta_field(seed=1)
for i in range(3):
    torch.cuda.synchronize(); t = time.perf_counter()
    e, pk, cnt = g.measure_power(nbins=32)
    torch.cuda.synchronize()
    print("measure_power %d^3: %.1f ms, %d modes" % (n, 1e3 * (time.perf_counter() - t), int(cnt.sum())))

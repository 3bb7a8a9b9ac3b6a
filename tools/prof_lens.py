"""Timing / profiling driver of the lensing-potential scan: prof_lens.py [n]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from randomfield_b200 import _lib, powertools
from randomfield_b200.generate import cot_k, lensing_tables
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
eng = _lib.Engine(n, n, n, dtype=np.float32)
k, sig = powertools.sigma_table(powertools.load_default_power(), (n, n, n), 2.0)
eng.set_power(k, sig, 2.0)
eng.generate_delta(1, True, np.full(n, -4.8e-36), False)
dc = np.arange(n) * 2.0
tabs = lensing_tables(dc, cot_k(dc, dc, 0.0))
s = torch.cuda.Stream(); eng.set_stream(s.cuda_stream)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(3):
    eng.lensing(n // 32, tabs)
    ev[0].record(s); eng.lensing(n // 32, tabs); ev[1].record(s); s.synchronize()
    print("lensing %d^3: %.3f ms (%.0f GB/s at 8 B/cell)" % (n, ev[0].elapsed_time(ev[1]), 8 * n ** 3 / ev[0].elapsed_time(ev[1]) / 1e6))

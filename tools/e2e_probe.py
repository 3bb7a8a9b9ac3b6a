import sys, time
sys.path.insert(0, ".")
import numpy as np, torch
import randomfield_b200 as rf
from randomfield_b200 import _lib
g = rf.Generator(1024, 1024, 1024, 2.0)
d = g.generate_delta_field(seed=1, save_potential=False)
e = g._engine
def t(f, n=3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(n): f(i)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("generate_delta_field      %.2f ms" % t(lambda i: g.generate_delta_field(seed=10 + i, save_potential=False)))
print("copy_field (pinned)       %.2f ms" % t(lambda i: e.copy_field(g._host.array)))
print("copy_rows whole (pinned)  %.2f ms" % t(lambda i: e.copy_rows(g._host.array, 0)))
def gen_only(i):
    e.generate_delta(20 + i, want_stats=True); e.sync()
print("generate_delta device     %.2f ms" % t(gen_only))
def streamed(i):
    e.generate_xpass(30 + i); e.finish_yz_streamed(g._host.array, nranges=8); e.copy_wait()
print("xpass+finish_streamed(8)  %.2f ms" % t(streamed))
def streamed16(i):
    e.generate_xpass(30 + i); e.finish_yz_streamed(g._host.array, nranges=32); e.copy_wait()
print("xpass+finish_streamed(32) %.2f ms" % t(streamed16))
def unstreamed(i):
    e.generate_delta(40 + i, want_stats=True); e.copy_field(g._host.array)
print("generate + copy_field     %.2f ms" % t(unstreamed))

"""Multi-GPU timing probe of the chunked exchange (run under torchrun): times the copy-engine
pushes alone, the kernels alone and the pipelined step.  usage: exchange_probe.py nx,ny,nz chunks"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from randomfield_b200 import powertools
from randomfield_b200.slab import SlabEngine

shape = tuple(int(v) for v in sys.argv[1].split(","))
chunks = int(sys.argv[2])
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
os.environ["RF_EXCHANGE"] = "chunked"
os.environ["RF_CHUNKS"] = str(chunks)
se = SlabEngine(*shape, 2.0, powertools.load_default_power())
eng, q = se.engine, se._comm
rank = dist.get_rank()


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def copies_only():
    eng.chunk_begin_step()
    for c in range(se.nchunks):
        eng.chunk_exchange(c)
    with torch.cuda.stream(q):
        dist.all_reduce(se._fence)


def kernels_only():
    for c in range(se.nchunks):
        eng.chunk_xpass(1, False, c)
    for c in range(se.nchunks):
        eng.chunk_arrived(c)
        eng.chunk_pass2(c)
    eng.finish_yz(None, True)


def full():
    se.generate(7)


for name, fn in (("copies", copies_only), ("kernels", kernels_only), ("pipelined", full)):
    ms = timed(fn)
    if rank == 0:
        print("%s chunks=%d: %s %.3f ms" % ("x".join(map(str, shape)), se.nchunks, name, ms), flush=True)
if rank == 0:
    print("bytes sent per rank %.3f GB" % (se.exchange_bytes / 1e9))
dist.destroy_process_group()

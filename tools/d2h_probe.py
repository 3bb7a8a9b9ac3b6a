"""Device-to-host copy rate per GPU when 1 .. N ranks copy at once (torchrun, one rank per GPU).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/d2h_probe.py

Each rank copies 4 GB from its GPU into two page-locked 512 MB stages (the pattern of
SlabEngine.iter_field_rows).  Phases: every subset size k = 1, 2, 4, .. N of ranks copying at the
same time (the others idle at the barrier), then the host's view of the box (NUMA nodes, cpus).
"""
import os
import sys
import time

import torch
import torch.distributed as dist

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

CHUNK = 512 << 20
TOTAL = 4 << 30
src = torch.empty(TOTAL, dtype=torch.uint8, device="cuda")
src.fill_(1)
stages = [torch.empty(CHUNK, dtype=torch.uint8).pin_memory() for _ in range(2)]
st = torch.cuda.Stream()


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def copy_all(chunk=CHUNK):
    ev = [None, None]
    with torch.cuda.stream(st):
        for i, off in enumerate(range(0, TOTAL, chunk)):
            b = i & 1
            if ev[b] is not None:
                ev[b].synchronize()
            stages[b][:chunk].copy_(src[off:off + chunk], non_blocking=True)
            ev[b] = torch.cuda.Event()
            ev[b].record(st)
    st.synchronize()


copy_all()
k = 1
while k <= world:
    for chunk in (CHUNK, CHUNK // 8):
        barrier()
        t0 = time.perf_counter()
        if rank < k:
            copy_all(chunk)
        dt = time.perf_counter() - t0
        barrier()
        rate = torch.tensor([TOTAL / dt / 1e9 if rank < k else 0.0], device="cuda")
        if world > 1:
            rates = [torch.zeros_like(rate) for _ in range(world)]
            dist.all_gather(rates, rate)
        else:
            rates = [rate]
        if rank == 0:
            rs = [float(r.item()) for r in rates[:k]]
            print("ranks copying: %d  chunk %4d MB  GB/s per rank: %s  sum %.1f" % (
                k, chunk >> 20, " ".join("%.1f" % r for r in rs), sum(rs)), flush=True)
    k *= 2

if rank == 0:
    try:
        nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))
    except OSError:
        nodes = []
    print("numa nodes:", nodes, " cpus:", os.cpu_count(), " affinity:", len(os.sched_getaffinity(0)))
    for line in open("/proc/meminfo").read().splitlines()[:3]:
        print(line)
if world > 1:
    dist.destroy_process_group()

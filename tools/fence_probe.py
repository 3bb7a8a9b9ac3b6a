"""Cost of the 1-element NCCL all-reduce the chunked exchange uses as its cross-rank fence.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/fence_probe.py"""
import os
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
x = torch.zeros(1, device="cuda")
for _ in range(20):
    dist.all_reduce(x)
torch.cuda.synchronize()
dist.barrier()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 500
a.record()
for _ in range(n):
    dist.all_reduce(x)
b.record()
torch.cuda.synchronize()
if dist.get_rank() == 0:
    print("world %d: %.2f us per 1-element all-reduce (back to back on one stream)" % (
        dist.get_world_size(), a.elapsed_time(b) * 1e3 / n), flush=True)
dist.destroy_process_group()

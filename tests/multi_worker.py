"""torchrun worker for tests/test_gpu_multi.py: one rank per GPU, NCCL all-to-all."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

from randomfield_b200 import powertools
from randomfield_b200.slab import SlabEngine

out = sys.argv[1]
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = SlabEngine(256, 256, 128, 2.0, powertools.load_default_power())
eng.generate(41)   # two steps: the second one reuses every buffer of the first
eng.generate(42)
mean, std = eng.global_stats()
np.save(os.path.join(out, "rank%d.npy" % dist.get_rank()), eng.local_field())
if dist.get_rank() == 0:
    open(os.path.join(out, "std.txt"), "w").write(repr(std))
    open(os.path.join(out, "mode.txt"), "w").write(eng.mode)
dist.barrier()
dist.destroy_process_group()

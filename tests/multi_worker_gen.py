"""torchrun worker for tests/test_gpu_multi.py::test_generator_over_two_ranks: the public Generator
API slab-decomposed over the ranks of the default process group (one rank per GPU)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

import randomfield_b200 as rf

out = sys.argv[1]
shape = tuple(int(v) for v in sys.argv[2].split(","))
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank = dist.get_rank()
gen = rf.Generator(*shape, 2.0, group=True)
res = {}
res["delta"] = gen.generate_delta_field(seed=77, smoothing_length_Mpc_h=1.0).copy()
res["rms"] = np.array(gen.delta_field_rms)
res["rho"] = gen.convert_delta_to_density().copy()
res["phi"] = gen.calculate_newtonian_potential(light_cone=True).copy()
res["psi"] = gen.calculate_lensing_potential(i_min=3).copy()
res["phi_after"] = gen.device_field().cpu().numpy()
full = gen.gather(res["phi"])
if rank == 0:
    res["phi_gathered"] = full
res["x_slab"] = np.array(gen.x_slab)
np.savez(os.path.join(out, "gen_rank%d.npz" % rank), **res)
dist.barrier()
dist.destroy_process_group()

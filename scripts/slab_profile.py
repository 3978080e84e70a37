"""Section profile of the slab step (host wall time with a device sync after every section; diagnostic only).
torchrun --nproc-per-node N scripts/slab_profile.py [n_side]"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_fluid_simulation_b200.slab import DistDriver, SlabRank, band_of_uniform_box  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
settings, ts, band = band_of_uniform_box(n_side, rank, world)
per_row = 4 * (int(np.ceil(settings.size[0] / settings.smoothing_radius)) + 2)
r = SlabRank(lr, settings, rank, world, capacity=int(band.shape[0] * 1.1) + 8 * per_row, pad=max(4 * per_row, 1 << 14),
             emig_cap=max(1 << 16, 8 * per_row))
r.upload_band(band)
drv = DistDriver(r)
for _ in range(3):
    drv.tick(ts)
drv.profile = {}
K = 10
for _ in range(K):
    drv.tick(ts)
prof = {k: round(1e3 * v / K, 4) for k, v in drv.profile.items()}
if rank == 0:
    print(json.dumps({"world": world, "n_side": n_side, "ms_per_section": prof, "sum_ms": round(sum(prof.values()), 3)}))
r.close()
dist.barrier()
dist.destroy_process_group()

#!/usr/bin/env python
"""torchrun --nproc-per-node N scripts/multi_parity.py [--n-side 2000] [--ticks 3] [--dam]: the multi-GPU step, one rank
per GPU over NCCL, gathered on rank 0 and compared bit for bit with a single-GPU run of the same array (the body of
bench.py's parity pre-check, gpu_fluid_simulation_b200/multi_bench.py)."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_fluid_simulation_b200 import multi_bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n-side", type=int, default=2000)
ap.add_argument("--ticks", type=int, default=3)
ap.add_argument("--dam", action="store_true")
ap.add_argument("--math", default="exact")
a = ap.parse_args()
rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local_rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
kind, arg = ("dam", (a.n_side // 4, a.n_side)) if a.dam else ("uniform", a.n_side)
res, bad = multi_bench.parity_vs_single_gpu(dist, rank, world, local_rank, kind, arg, ticks=a.ticks, math_mode=1 if a.math == "fast" else 0,
                                            **({"window_near": 65536} if a.dam else {}))
if rank == 0:
    print(json.dumps(res), flush=True)
    print(f"world {world} fields_differing {res['fields_differing']} transport {res['transport']}", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if bad == 0 else 1)

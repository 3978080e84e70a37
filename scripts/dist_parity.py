"""torchrun --nproc-per-node N scripts/dist_parity.py : slab mode over NCCL vs a single-GPU run of the same state.
Every rank runs its band through DistDriver; rank 0 gathers the bands, runs the whole scene on its own GPU and
requires bit-identical keys, start_indices, density, velocity and position."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import scenes, slab

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ok = True
cases = [(n, sc, t, res) for (n, sc, t) in (("U400", scenes.uniform_box(400), 12), ("D200x800", scenes.dam_break(200, 800), 25))
         for res in (True, False)]
for name, (s, ts, p), ticks, resolved in cases:
    name = f"{name}/{'device' if resolved else 'host'}-resolved"
    r = slab.SlabRank(local, s, rank, world, capacity=p.shape[0], pad=1 << 14)
    bands = slab.split_particles(p, s, r.bounds)
    r.upload_band(bands[rank])
    drv = slab.DistDriver(r, device_resolved=resolved)
    for _ in range(ticks):
        drv.tick(ts)
    r.sync()
    mine = r.download_band()
    lo, hi = r.owns_rows()
    table = r.sim.download_start_indices()
    gathered = [None] * world if rank == 0 else None
    dist.gather_object((mine, table, lo, hi), gathered, dst=0)
    if rank == 0:
        order = np.concatenate(bands)
        with fluid.FluidSimulation.new(local, s) as ref:
            ref.upload_particles(order)
            for _ in range(ticks):
                ref.tick(ts)
            want, want_table, gw = ref.download_particles(), ref.download_start_indices(), ref.grid_w
        got = np.concatenate([g[0] for g in gathered])
        tab = np.zeros_like(want_table)
        for g in gathered:
            tab[g[2] * gw:g[3] * gw] = g[1][g[2] * gw:g[3] * gw]
        same = got.shape == want.shape and np.array_equal(tab, want_table) and all(np.array_equal(got[f], want[f]) for f in got.dtype.names)
        print(f"dist parity {name} world={world} ticks={ticks} counts={[g[0].shape[0] for g in gathered]}: {'BIT-IDENTICAL' if same else 'MISMATCH'}", flush=True)
        ok = ok and same
    r.close()
    dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)

"""Small end-to-end exercise of every kernel of the tick, with cross-checks between the sort modes (written for a
compute-sanitizer memcheck pass; the tool is closed on this pool, so it runs plain: python scripts/sanitize_smoke.py).
Covers the one-launch and the two-kernel cell scan, all three sort modes, ragged counts, the pipelined transfers,
hash-sort-only ticks and a 3-band slab run on one GPU."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes, slab  # noqa: E402


def run(s, ts, p, mode, ticks=2):
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.sort_mode = mode
        sim.upload_particles(p)
        for _ in range(ticks):
            sim.tick(ts)
        sim.tick_hash_sort_only(ts)
        return sim.download_particles(), sim.download_permutation(), sim.download_start_indices()


for name, scene in (("small grid", scenes.uniform_box(97)), ("large grid", scenes.dam_break(300, 700)),
                    ("ragged", scenes.uniform_box(31))):
    s, ts, p = scene
    outs = [run(s, ts, p, m) for m in (0, 1, 2)]
    for o in outs[1:]:
        for a, b in zip(outs[0], o):
            assert np.array_equal(a, b), name
    print("ok", name, s.particle_count, "particles", flush=True)

s, ts, p = scenes.uniform_box(64)
n = p.shape[0]
buf = torch.from_numpy(p.view(np.uint8).reshape(-1).copy()).pin_memory()
out = torch.zeros(n * 32, dtype=torch.uint8).pin_memory()
with fluid.FluidSimulation.new(0, s) as sim:
    for _ in range(3):
        sim.upload_particles_async(buf)
        sim.tick(ts)
        sim.download_particles_async(out)
    sim.sync_transfers()
print("ok pipelined transfers", flush=True)

s, ts, p = scenes.uniform_box(96)
ranks = [slab.SlabRank(0, s, r, 3, capacity=p.shape[0], pad=4096) for r in range(3)]
bands = slab.split_particles(p, s, ranks[0].bounds)
for r, b in zip(ranks, bands):
    r.upload_band(b)
drv = slab.LocalDriver(ranks)
for _ in range(4):
    drv.tick(ts)
got = np.concatenate([r.download_band() for r in ranks])
paths = [r.sim.last_sort_path for r in ranks]
for r in ranks:
    r.close()
assert got.shape[0] == p.shape[0] and all(x == 2 for x in paths), paths
print("ok slab, 3 bands, counting path", flush=True)

#!/usr/bin/env python
"""Where does the end-to-end loop of bench.py (upload_async + tick + download_async per step) spend its time?
Host timestamps per call, for sort mode AUTO (waits for the scan's word after an upload) and COUNTING (no wait)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import scenes  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
s, ts, p = scenes.uniform_box(side)
n = s.particle_count
with fluid.FluidSimulation.new(0, s) as sim:
    sim.upload_particles(p)
    sim.tick_n(ts, 2)
    host_in = torch.empty((n, 8), dtype=torch.float32).pin_memory()
    outs = [torch.empty((n, 8), dtype=torch.float32).pin_memory() for _ in range(2)]
    views = [o.numpy().view(fluid.PARTICLE_DTYPE).reshape(-1) for o in outs]
    host_in.numpy().view(np.uint8).reshape(-1)[:] = sim.download_particles().view(np.uint8).reshape(-1)
    for mode in (0, 2, 0, 2):
        sim.sort_mode = mode
        sim.upload_particles_async(host_in); sim.tick(ts); sim.download_particles_async(views[0]); sim.sync_transfers()
        k = 6
        marks = []
        t0 = time.perf_counter()
        for j in range(k):
            a = time.perf_counter(); sim.upload_particles_async(host_in)
            b = time.perf_counter(); sim.tick(ts)
            c = time.perf_counter(); sim.download_particles_async(views[j & 1])
            d = time.perf_counter()
            marks.append((1e3 * (a - t0), 1e3 * (b - a), 1e3 * (c - b), 1e3 * (d - c)))
        sim.sync_transfers()
        dt = time.perf_counter() - t0
        print(f"sort_mode {mode}: {1e3 * dt / k:.2f} ms per step; per step (start, upload_async, tick, download_async) ms:",
              [tuple(round(x, 2) for x in m) for m in marks], flush=True)

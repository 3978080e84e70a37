#!/usr/bin/env python
"""Where does the late dam break's time go?  Runs 1M-D for --ticks ticks, then reports per-phase
times next to the cell statistics that drive the neighbour passes' work: candidates per particle (sum of the populations of
the 3x3 cells), neighbours per particle, and the share of 256-particle tiles whose three row spans fit the staging buffer."""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--ticks", type=int, nargs="+", default=[20, 320])
ap.add_argument("--cap", type=int, default=1920)
a = ap.parse_args()
s, ts, p = scenes.dam_break(500, 2000)
out = []
with fluid.FluidSimulation.new(0, s) as sim:
    sim.upload_particles(p)
    done = 0
    for target in a.ticks:
        sim.tick_n(ts, target - done)
        done = target
        sim.sync()
        sim.enable_timing(True)
        sim.tick_n(ts, 50)
        done += 50
        phases = dict(sim.last_timing())
        sim.enable_timing(False)
        q = sim.download_particles()
        gw, gh = sim.grid_w, sim.grid_h
        keys = q["grid"].astype(np.int64)
        pop = np.bincount(keys, minlength=gw * gh).reshape(gh, gw)
        # candidates per particle = populations of the 3x3 cells around its cell
        pad = np.pad(pop, 1)
        box = sum(pad[1 + dy:1 + dy + gh, 1 + dx:1 + dx + gw] for dy in (-1, 0, 1) for dx in (-1, 0, 1))
        cand = box.reshape(-1)[keys]
        # tiles of 256 sorted particles: rows y-1..y+1 x cells [x_lo-1, x_hi+1] per row group (kernels_sph.cuh make_tile_plan)
        csum = np.concatenate([[0], np.cumsum(pop.reshape(-1))])
        n = keys.shape[0]
        fits = 0
        tiles = (n + 255) // 256
        for t in range(tiles):
            k0, k1 = keys[t * 256], keys[min(n, t * 256 + 256) - 1]
            y0, x0, y1, x1 = k0 // gw, k0 % gw, k1 // gw, k1 % gw
            total = 0
            if y1 - y0 + 1 <= 2:
                for g, y in enumerate(range(y0, y1 + 1)):
                    xl = x0 if g == 0 else 0
                    xh = x1 if y == y1 else gw - 1
                    for r in (y - 1, y, y + 1):
                        if 0 <= r < gh:
                            total += csum[r * gw + min(xh + 2, gw)] - csum[r * gw + max(xl - 1, 0)]
            else:
                total = csum[min((y1 + 1) * gw + min(x1 + 2, gw), gw * gh)] - csum[max((y0 - 1) * gw + max(x0 - 1, 0), 0)]
            fits += total <= a.cap
        rec = {"after_ticks": done, "phases_ms": {k: round(v, 4) for k, v in phases.items()}, "ms_per_step": round(sum(v for k, v in phases.items()), 4),
               "max_cell": int(pop.max()), "mean_pop_of_occupied": float(pop[pop > 0].mean()), "candidates_per_particle_mean": float(cand.mean()),
               "candidates_p99": float(np.percentile(cand, 99)), "tiles": int(tiles), "tiles_staged_frac": fits / tiles,
               "sort_path": "counting" if sim.last_sort_path == 2 else "radix"}
        out.append(rec)
        print(json.dumps(rec), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/crowding_probe.json", "w"), indent=1)

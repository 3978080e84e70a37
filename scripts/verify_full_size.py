#!/usr/bin/env python
"""Full-size parity: the CUDA path against the CPU oracle on the SAME 64M-particle scene bench.py times
(BASELINE.json config 5), every field of every particle compared bit for bit after each tick.

    python scripts/verify_full_size.py [n_side=8000] [ticks=2]

Prints one JSON line. The oracle is test infrastructure (see the header of oracle/oracle.py); this script is a checker, not product code.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import scenes  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    n_side = int(sys.argv[1]) if len(sys.argv) > 1 else 8000
    ticks = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    O.build()
    s, ts, p = scenes.uniform_box(n_side)
    osim = O.OracleSim(s, p, sort_mode=O.SORT_COUNTING)  # stable counting sort: same order as the stable radix sort
    out = {"particles": int(s.particle_count), "ticks": ticks, "fields": {}, "cpu_s_per_tick": [], "identical": True}
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        del p
        for t in range(ticks):
            sim.tick(ts)
            t0 = time.perf_counter()
            osim.tick(ts)
            out["cpu_s_per_tick"].append(round(time.perf_counter() - t0, 2))
            got = sim.download_particles()
            table = sim.download_start_indices()
            for f in ("grid", "position", "predicted_position", "velocity", "density"):
                a = np.ascontiguousarray(got[f]).view(np.uint32)
                b = np.ascontiguousarray(osim.p[f]).view(np.uint32)
                diff = int((a != b).sum())
                out["fields"][f"tick{t + 1}.{f}"] = diff
                out["identical"] &= diff == 0
            diff = int((table != osim.table).sum())
            out["fields"][f"tick{t + 1}.start_indices"] = diff
            out["identical"] &= diff == 0
            del got
        out["sort_path"] = {1: "radix", 2: "counting"}[sim.last_sort_path]
    print(json.dumps(out), flush=True)
    return 0 if out["identical"] else 1


if __name__ == "__main__":
    sys.exit(main())

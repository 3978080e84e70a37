#!/usr/bin/env python
"""Same-box A/B of two builds of libfluid_b200.so (FLUID_B200_LIB selects the library): per-phase times of a few scenes.
usage: FLUID_B200_LIB=/path/to/lib.so python scripts/ab_quick.py [64m]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import scenes  # noqa: E402


def run(name, s, ts, p, warm, steps):
    with fluid.FluidSimulation.new(0, s) as sim:
        if p is not None:
            sim.upload_particles(p)
        sim.tick_n(ts, warm)
        sim.sync()
        sim.enable_timing(True)
        sim.tick_n(ts, steps)
        ph = dict(sim.last_timing())
        sim.enable_timing(False)
    print(json.dumps({"scene": name, "ms": round(sum(ph.values()), 4), "phases": {k: round(v, 4) for k, v in ph.items()}}), flush=True)


s, ts, p = scenes.uniform_box(1000)
run("1m-u", s, ts, p, 20, 200)
s, ts, p = scenes.dam_break(500, 2000)
run("1m-d", s, ts, p, 20, 200)
run("1m-d@320", s, ts, p, 320, 100)
run("1m-d@470", s, ts, p, 470, 100)
if "64m" in sys.argv:
    s, ts, p = scenes.uniform_box(8000)
    run("64m-u", s, ts, p, 5, 20)

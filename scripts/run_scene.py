#!/usr/bin/env python
"""Run one scene for --ticks ticks on GPU 0 and exit (a target for ncu: no timing, no downloads inside the run).
usage: python scripts/run_scene.py --workload 1m-d --ticks 325"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_fluid_simulation_b200 as fluid  # noqa: E402
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="1m-d", choices=["1m-u", "1m-d", "16m-u", "64m-u", "default"])
ap.add_argument("--ticks", type=int, default=8)
ap.add_argument("--one-call", action="store_true", help="fluid_sim_tick_n instead of one call per tick")
a = ap.parse_args()
if a.workload == "default":
    s, ts, p = SimulationSettings(), TickSettings(), None
elif a.workload == "1m-d":
    s, ts, p = scenes.dam_break(500, 2000)
else:
    s, ts, p = scenes.uniform_box({"1m-u": 1000, "16m-u": 4000, "64m-u": 8000}[a.workload])
with fluid.FluidSimulation.new(0, s) as sim:
    if p is not None:
        sim.upload_particles(p)
    if a.one_call:
        sim.tick_n(ts, a.ticks)
    else:
        for _ in range(a.ticks):
            sim.tick(ts)
    sim.sync()
    print("ticks", sim.tick_count, "particles", sim.particle_count, "sort path", sim.last_sort_path)

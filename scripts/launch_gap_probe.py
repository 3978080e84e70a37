"""How much of a small scene's step is launch gaps: 1M-U ticks timed with CUDA events around tick_n, with and without the
per-phase event marks, against the sum of the per-phase times. Run on a GPU box: python scripts/launch_gap_probe.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import scenes

out = {}
for name, n_side in (("1m-u", 1000), ("250k-u", 500)):
    s, ts, p = scenes.uniform_box(n_side)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        stream = torch.cuda.ExternalStream(sim.stream, device=0)
        sim.tick_n(ts, 20)
        sim.sync()
        res = {}
        for marks in (False, True):
            sim.enable_timing(marks)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(stream):
                e0.record(stream)
                sim.tick_n(ts, 200)
                e1.record(stream)
            sim.sync()
            res["marks" if marks else "plain"] = e0.elapsed_time(e1) / 200
            if marks:
                res["phases_ms"] = dict(sim.last_timing())
        sim.enable_timing(False)
        out[name] = res
print(json.dumps(out))

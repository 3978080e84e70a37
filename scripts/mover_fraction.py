import sys, numpy as np
sys.path.insert(0, '/root/repo')
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import scenes
for name, (s, ts, p) in {"16m-u": scenes.uniform_box(4000), "1m-d": scenes.dam_break(500, 2000)}.items():
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        out = []
        for t in range(1, 41):
            prev = None
            if t in (2, 5, 10, 20, 40):
                prev = sim.download_particles()["grid"].copy()
            sim.tick(ts)
            if prev is not None:
                ku = sim.download_unsorted_keys()
                d = ku.astype(np.int64) - prev.astype(np.int64)
                out.append((t, float((d != 0).mean()), int(np.abs(d).max())))
        print(name, out, flush=True)

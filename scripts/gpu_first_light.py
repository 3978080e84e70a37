"""First-light check on the GPU box: one tick of a few scenes against the oracle, timing of the 1M scenes."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import scenes
from oracle import oracle as O
import parity

def check(name, settings, ts, p, n_ticks=1):
    t0 = time.time()
    osim, sim, extras = parity.run_pair(fluid, O, settings, ts, p, n_ticks=n_ticks)
    try:
        stats = parity.compare_state(osim, sim, extras, ts)
        print(name, "PARITY OK", json.dumps(stats), f"{time.time()-t0:.1f}s", flush=True)
    except AssertionError as e:
        print(name, "PARITY FAIL", e, flush=True)
    sim.close()

s, ts, p = scenes.uniform_box(64)
check("U64", s, ts, p)
s, ts, p = scenes.uniform_box(256)
check("U256", s, ts, p)
check("U256x5", s, ts, p, n_ticks=5)
s, ts, p = scenes.dam_break(64, 256)
check("D64x256", s, ts, p)
check("D64x256x20", s, ts, p, n_ticks=20)
s = fluid.SimulationSettings()
check("default", s, fluid.TickSettings(), O.init_lattice(s))

for name, (s, ts, p) in {"1M-U": scenes.uniform_box(1000), "1M-D": scenes.dam_break(500, 2000), "4M-U": scenes.uniform_box(2000), "16M-U": scenes.uniform_box(4000)}.items():
    sim = fluid.FluidSimulation.new(0, s)
    sim.upload_particles(p)
    sim.tick_n(ts, 10); sim.sync()
    t0 = time.time(); sim.tick_n(ts, 50); sim.sync(); dt = (time.time() - t0) / 50
    sim.enable_timing(True); sim.tick(ts); tm = sim.last_timing(); sim.enable_timing(False)
    n = s.particle_count
    print(name, f"{dt*1e3:.3f} ms/tick  {n/dt:.3e} particle-steps/s  200B-roofline {200*n/dt/1e9:.0f} GB/s", [(a, round(b, 4)) for a, b in tm], flush=True)
    sim.close()

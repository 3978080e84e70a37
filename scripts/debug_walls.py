import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings
from oracle import oracle as O
import parity
n = 2048
s = SimulationSettings(particle_count=n, size=(8.0, 8.0), texture_size=(8, 8))
rng = np.random.default_rng(3)
p = np.zeros(n, dtype=O.PARTICLE_DTYPE)
p["position"] = (rng.random((n, 2)) - 0.5) * 8.0
p["velocity"] = rng.normal(size=(n, 2)) * 400
if "--nan" in sys.argv:
    p["velocity"][5] = [np.nan, 1.0]
    p["velocity"][6] = [np.inf, 0.0]
p["position"][7] = [4.0, -4.0]
p["predicted_position"] = p["position"]
for ticks in (1, 2):
    osim, sim, extras = parity.run_pair(fluid, O, s, TickSettings(gravity=(0.0, 9.8)), p, n_ticks=ticks, debug_force=False)
    got = sim.download_particles()
    print("ticks", ticks, "keys eq", np.array_equal(got["grid"], osim.p["grid"]), "table eq", np.array_equal(sim.download_start_indices(), osim.table))
    for f in ("predicted_position", "density", "velocity", "position"):
        a, b = got[f], osim.p[f]
        bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
        if bad.ndim > 1: bad = bad.any(axis=1)
        print("  ", f, "mismatches", int(bad.sum()), "first", np.nonzero(bad)[0][:5])
        if bad.any():
            i = np.nonzero(bad)[0][0]
            print("     got", got[i], "\n     ref", osim.p[i])
    sim.close()

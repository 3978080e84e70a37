"""BASELINE.md §3 runs B1-B5: the CPU restatement (oracle, test infrastructure) timed on this machine's host cores.
Prints one JSON line; `nproc` is recorded. Run anywhere (no GPU): python scripts/cpu_baselines.py"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import oracle as O  # noqa: E402
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes  # noqa: E402

out = {"nproc": os.cpu_count(), "affinity": len(os.sched_getaffinity(0))}
O.lib().oracle_set_num_threads(len(os.sched_getaffinity(0)))


def timed_ticks(sim, ts, ticks, **kw):
    t0 = time.perf_counter()
    for _ in range(ticks):
        sim.tick(ts, **kw)
    return (time.perf_counter() - t0) / ticks


s, ts = SimulationSettings(), TickSettings()
for name, mode in (("stable_sort", O.SORT_STABLE), ("bitonic_schedule", O.SORT_BITONIC)):
    sim = O.OracleSim(s, O.init_lattice(s), sort_mode=mode)
    per = timed_ticks(sim, ts, 100 if mode == O.SORT_STABLE else 20)
    out[f"B1_default_scene_{name}"] = {"s_per_tick": round(per, 5), "particle_steps_per_s": round(s.particle_count / per)}

s, ts, p = scenes.dam_break(500, 2000)
sim = O.OracleSim(s, p, sort_mode=O.SORT_COUNTING)
per = timed_ticks(sim, ts, 20)
out["B2_1m_dam_break_20_ticks"] = {"s_per_tick": round(per, 4), "particle_steps_per_s": round(s.particle_count / per)}

s, ts, p = scenes.uniform_box(1000)
sim = O.OracleSim(s, p, sort_mode=O.SORT_COUNTING)
per = timed_ticks(sim, ts, 5, stop_after=1)
out["B3_1m_uniform_hash_sort_table"] = {"s_per_step": round(per, 4)}

# the reference's own shaders (transpiled WGSL, oracle/_ref) with their own dispatch schedule, where the library is there
from oracle import wgsl_ref as W  # noqa: E402
if W.available():
    W.lib().wgslref_set_num_threads(len(os.sched_getaffinity(0)))
    s, ts = SimulationSettings(), TickSettings()
    sim = W.WgslRefSim(O, s)
    per = timed_ticks(sim, ts, 20)
    out["R1_default_scene_transpiled_wgsl"] = {"s_per_tick": round(per, 5), "particle_steps_per_s": round(s.particle_count / per)}
    s, ts, p = scenes.dam_break(500, 2000)
    sim = W.WgslRefSim(O, s, p)
    per = timed_ticks(sim, ts, 5)
    out["R2_1m_dam_break_transpiled_wgsl"] = {"s_per_tick": round(per, 4), "particle_steps_per_s": round(s.particle_count / per)}

for size in (1024, 4096, 8192):
    rng = np.random.default_rng(size)
    mask = (rng.random((size, size)) < 0.002).astype(np.uint8) * 255
    if hasattr(O, "jfa_seed_from_mask"):
        seeds = O.jfa_seed_from_mask(mask)
        t0 = time.perf_counter()
        O.jfa_run(seeds)
        out[f"B4_jfa_oracle_{size}"] = {"ms_per_field": round(1e3 * (time.perf_counter() - t0), 1)}
    if hasattr(O, "generate_smooth_gradient_field") and size == 1024:
        t0 = time.perf_counter()
        O.generate_smooth_gradient_field(mask)
        out["B5_cpu_field_1024_single_thread"] = {"ms_per_field": round(1e3 * (time.perf_counter() - t0), 1)}
print(json.dumps(out))
if os.path.isdir("gpurun_out"):
    json.dump(out, open("gpurun_out/cpu_baselines.json", "w"), indent=1)

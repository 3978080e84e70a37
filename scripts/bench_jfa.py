"""BASELINE.json config 4: jump-flood field at 4096^2 and 8192^2 from 4M particles on one B200.
Reports ms per full field (device time of the JFA passes only, library CUDA events), pixel-passes/s, achieved GB/s
against the algorithmic 16 B/pixel/pass of the reference's rg32uint format (the library moves 8 B/pixel/pass: packed
ushort2), and the CPU oracle beside it. Exactness is checked against the oracle at 4096^2."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import gpu_fluid_simulation_b200 as fluid
from gpu_fluid_simulation_b200 import scenes
from oracle import oracle as O

peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
out = {"peak_gbs": peak, "runs": []}
for scene_name, (s, ts, p) in (("4M-D", scenes.dam_break(1000, 4000)), ("4M-U", scenes.uniform_box(2000))):
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        for size in (4096, 8192):
            passes = int(np.ceil(np.log2(size)))
            for _ in range(3):
                sim.jfa_from_particles(size, size, want_seeds=False, want_field=False)
            ms = [sim.jfa_from_particles(size, size, want_seeds=False, want_field=False)[2] for _ in range(10)]
            t = float(np.median(ms))
            rec = {"scene": scene_name, "size": size, "passes": passes, "ms_per_field": t,
                   "pixel_passes_per_s": size * size * passes / (t * 1e-3),
                   "achieved_gbs_algorithmic_16B": 16.0 * size * size * passes / (t * 1e-3) / 1e9,
                   "frac_of_measured_peak_16B": 16.0 * size * size * passes / (t * 1e-3) / 1e9 / peak,
                   "moved_gbs_packed_8B_ideal": 8.0 * size * size * passes / (t * 1e-3) / 1e9}
            if size == 4096:
                seeds, fld, _ = sim.jfa_from_particles(size, size)
                t0 = time.perf_counter()
                ref = O.jfa_run(O.jfa_seed_from_particles(p, s.size, size, size))
                rec["cpu_oracle_ms"] = 1e3 * (time.perf_counter() - t0)
                rec["cpu_threads"] = int(O.lib().oracle_num_threads())
                rec["exact_vs_oracle"] = bool(np.array_equal(seeds, ref))
            out["runs"].append(rec)
            print(json.dumps(rec), flush=True)
# the reference's own CPU routine at its own size (src/renderer.rs:16: 1024^2)
mask = np.full((1024, 1024), 255, np.uint8); mask[300:700, 200:500] = 0
t0 = time.perf_counter(); O.generate_smooth_gradient_field(mask); out["generate_smooth_gradient_field_1024_ms"] = 1e3 * (time.perf_counter() - t0)
t0 = time.perf_counter(); fluid.jfa.to_force_field(fluid.jfa.jfa_run(fluid.jfa.seed_from_mask(mask))); out["gpu_mask_to_field_1024_ms_incl_host_copies"] = 1e3 * (time.perf_counter() - t0)
print(json.dumps({k: v for k, v in out.items() if k != "runs"}))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/jfa_bench.json", "w"), indent=1)

#!/usr/bin/env python
"""bench.py — particle-steps/s of the full SPH fluid step (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload 64m-u|16m-u|4m-u|1m-u|1m-d] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

A "step" is one complete tick (predict, hash, sort, cell tables, density, force+integrate) of the workload.
Default workload: the 64M-particle uniform box of BASELINE.json config 5 (the size the metric's target is quoted
on; ~4 GB of particle state, far larger than the 126 MB L2, so no L2 flush is needed between steps). At N > 1 the
same 64M box is cut into bands of grid rows over the ranks (strong scaling); the step and its one NCCL exchange per tick run
inside the library (fluid_multi_tick), and the run first checks the NCCL path bit for bit against one GPU.

One JSON line is printed by rank 0. `value` is device-timed (CUDA events on the library's stream, inputs resident
in HBM); `e2e` is measured through the public API with HOST buffers (pinned upload of the full particle state +
tick + download of the full state every step; `e2e.frame_pattern` adds the reference's own per-frame pattern: state
resident, force-field upload + tick + headless render downloaded). `roofline` is for the dominant kernel (force+integrate) with its
duration measured live by the library's per-phase CUDA events inside the timed region. `cpu_baseline` times the
CPU oracle (the restatement of the reference's WGSL; the reference itself cannot run here) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {  # name -> (kind, args)
    "64m-u": ("uniform", 8000), "16m-u": ("uniform", 4000), "4m-u": ("uniform", 2000), "1m-u": ("uniform", 1000),
    "1m-d": ("dam", (500, 2000)), "4m-d": ("dam", (1000, 4000)), "default": ("default", None),
}
# algorithmic bytes per particle-step, by pass
ALGO_BYTES = {"predict+key+hist": 28, "sort": 52, "gather+cells": 52 + 8, "density": 16, "force+integrate": 44}
ALGO_BYTES_TOTAL = 200
# bytes the kernels of a phase actually read + write per particle, by sort path
MOVED_BYTES = {"counting": {"sort": 14, "gather+cells": 49, "density": 8 + 4 + 4 + 24, "force+integrate": 8 + 8 + 8 + 4 + 24 + 16 + 4},
               "radix": {"predict+key+hist": 20, "sort": 44, "gather+cells": 60, "density": 8 + 4 + 4 + 24, "force+integrate": 8 + 8 + 8 + 4 + 24 + 16 + 4}}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """samples nvidia-smi SM clocks + throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append((time.time(), l)) for l in self.proc.stdout], daemon=True)
            self.t.start()
            t_end = time.time() + 3.0  # nvidia-smi needs a few hundred ms to print its first line
            while not self.lines and time.time() < t_end:
                time.sleep(0.02)
        except Exception:
            self.proc = None

    def mark(self):
        """wall-clock mark: samples are attributed to the interval between two marks"""
        return time.time()

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, l in self.lines:
            if t0 is not None and not (t0 <= t <= t1 + 0.12):  # a -lms 100 sample describes the 100 ms before it
                continue
            parts = [x.strip() for x in l.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for nme, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_scene(name):
    from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes
    kind, arg = WORKLOADS[name]
    if kind == "uniform":
        return scenes.uniform_box(arg)
    if kind == "dam":
        return scenes.dam_break(*arg)
    return SimulationSettings(), TickSettings(), None  # default lattice is built by create()


def cpu_sample_scene(workload, full=False):
    """the CPU arm's scene: the named workload itself (full=True) or a bounded sample of it — same spacing / h / tick
    parameters, 1M particles."""
    from gpu_fluid_simulation_b200 import scenes
    kind, arg = WORKLOADS[workload]
    if full:
        return make_scene(workload), f"the {workload} workload itself"
    if kind == "dam":
        return scenes.dam_break(500, 2000), "1M-particle dam-break (same spacing, h and tick parameters)"
    if kind == "default":
        return make_scene(workload), "the reference's default scene (100,000 particles) itself"
    return scenes.uniform_box(1000), "1M-particle sub-box of the uniform box (same spacing, h, jitter and tick parameters)"


def host_cores():
    """the cores this process may run on (torchrun exports OMP_NUM_THREADS=1 to its workers: the CPU arm must not inherit
    that, it is the reference's all-core CPU figure)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def run_cpu(workload, ticks, warmup=1, budget_s=25.0, full=False, kind="auto"):
    """times the reference's step on every host core.
    kind "reference": the reference's own WGSL shaders (compute.wgsl + funcs.wgsl, sort.wgsl), transpiled mechanically to
      C++ (oracle/transpile_wgsl.py -> oracle/_ref/libwgsl_ref.so) and dispatched as src/simulation.rs:512-537 records them:
      S(S+1)/2 bitonic passes over whole records, the 7x7 density walk, two 3x3 walks in move_particle. OpenMP over invocations.
    kind "port": the hand-written oracle with a stable counting sort (a fair CPU algorithm, not the reference's).
    kind "auto": "reference" when the transpiled library is there, else "port"."""
    from oracle import oracle as O
    from oracle import wgsl_ref as W
    O.build()
    if kind == "auto":
        kind = "reference" if W.available() else "port"
    cores = host_cores()
    O.lib().oracle_set_num_threads(cores)
    (s, ts, p), desc = cpu_sample_scene(workload, full)
    if kind == "reference":
        W.lib().wgslref_set_num_threads(cores)
        sim = W.WgslRefSim(O, s, p)
        how = ("the reference's WGSL transpiled to C++ (oracle/_ref), its own dispatch schedule (bitonic network, 7x7 density walk), "
               f"OpenMP over invocations on {cores} threads")
    else:
        sim = O.OracleSim(s, p, sort_mode=O.SORT_COUNTING)
        how = f"oracle port, OpenMP over particles on {cores} threads, stable counting sort"
    del p
    t_w = time.perf_counter()
    warm_done = 0
    for _ in range(warmup):
        sim.tick(ts)
        warm_done += 1
        if budget_s and time.perf_counter() - t_w > 0.25 * budget_s:
            break
    t0 = time.perf_counter()
    done = 0
    while done < ticks:
        sim.tick(ts)
        done += 1
        if budget_s and time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    n = s.particle_count
    return {"value": n * done / dt, "unit": "particle-steps/s", "cores": cores, "kind": kind,
            "sample": f"{done} ticks (after {warm_done} warm-up) of {desc}; {how}", "ms_per_step": 1e3 * dt / done,
            "particles": n, "ticks": done, "warmup": warm_done}


def reference_arm(args, rank):
    """--impl reference: the reference's own step on the host cores. Its WGSL cannot run through a graphics API on this box;
    what runs is that same WGSL transpiled to C++ (kind "reference"; the hand-written port if the transpiled library is
    missing), on the NAMED workload (config.same_config), every step one full tick; steps / warm-up are cut short when the
    time budget runs out (a 64M-particle tick of the reference's 351-pass bitonic network takes tens of seconds on a CPU)."""
    if rank != 0:
        return
    full = os.environ.get("FLUID_REF_SAMPLE", "") != "1"
    r = run_cpu(args.workload, args.steps, warmup=args.warmup, budget_s=150.0, full=full)
    line = {"impl": "reference", "metric": "particle-steps/sec", "value": r["value"], "unit": "particle-steps/s",
            "n_gpus": args.gpus, "steps": r["ticks"], "warmup": r["warmup"], "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles": r["particles"], "same_config": bool(full),
                       "note": "the reference's step on the host cores (its wgpu/WGSL path through a graphics API is unavailable on this "
                               "box): " + r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "particle-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="64m-u", choices=sorted(WORKLOADS))
    ap.add_argument("--math", default="exact", choices=["exact", "fast"],
                    help="force-pass arithmetic: exact = IEEE-rounded reference order (bit-identical to the oracle); "
                         "fast = hardware rsqrt/rcp + FMA (<= 1e-5 parity, tests/test_gpu_parity.py)")
    ap.add_argument("--sort", default="auto", choices=["auto", "radix", "counting"],
                    help="key-order path of the single-GPU tick (include/fluid_b200.h FLUID_SORT_*): identical results")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the short 1M / default-scene / fast-math timings appended to the line")
    ap.add_argument("--no-parity", action="store_true", help="--gpus > 1: skip the NCCL-path-vs-single-GPU parity check before timing")
    ap.add_argument("--e2e-steps", type=int, default=40,
                    help="steps of the end-to-end loop; its timed region includes filling and draining the transfer pipeline (one full "
                         "download after the last tick), which 6 steps did not amortise: 48.7 ms per step at 64M against a cadence of 42.3")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    if world != args.gpus:
        if args.gpus == 1 and world == 1:
            pass
        else:
            raise SystemExit(f"--gpus {args.gpus} needs WORLD_SIZE={args.gpus} (launch with torch.distributed.run); got {world}")

    import torch
    import gpu_fluid_simulation_b200 as fluid
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the fluid step has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        from gpu_fluid_simulation_b200.multi_bench import run_multi_bench
        line = run_multi_bench(args, rank, world, local_rank, WORKLOADS, ALGO_BYTES, *peaks(), ClockSampler)
        if rank == 0:
            print(json.dumps(line), flush=True)
        return

    peak, peak_src = peaks()
    s, ts, p = make_scene(args.workload)
    n = s.particle_count
    sim = fluid.FluidSimulation.new(local_rank, s)
    sim.math_mode = 1 if args.math == "fast" else 0
    sim.sort_mode = {"auto": 0, "radix": 1, "counting": 2}[args.sort]
    if p is not None:
        sim.upload_particles(p)
    sim.sync()
    stream = torch.cuda.ExternalStream(sim.stream, device=local_rank)

    # ---- device-resident throughput -------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    sim.tick_n(ts, args.warmup)
    sim.sync()
    launches0 = sim.launch_count
    # per-phase event marks (6 cudaEventRecord per tick) cost ~20 us per tick (profiles/r1d_launch_gap_probe.json):
    # noise at 64M (0.3 %), 12 % at 1M. Small scenes therefore time `value` without the marks and take the phase times
    # from a second pass of the same K steps right after the timed region.
    marks_in_region = n >= 8_000_000
    sim.enable_timing(marks_in_region)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t_clk0 = sampler.mark()
    with torch.cuda.stream(stream):
        e0.record(stream)
        sim.tick_n(ts, args.steps)
        e1.record(stream)
    sim.sync()
    torch.cuda.synchronize()
    t_clk1 = sampler.mark()
    ms = e0.elapsed_time(e1)
    launches_timed = sim.launch_count - launches0
    if not marks_in_region:
        sim.enable_timing(True)
        sim.tick_n(ts, args.steps)
        sim.sync()
    phases = sim.last_timing()
    between_ticks_ms = dict(phases).get("between ticks")  # device idle between two ticks (0 while the host keeps up)
    phases = [(k_, v) for k_, v in phases if k_ in ALGO_BYTES]
    sort_path = "counting" if sim.last_sort_path == 2 else "radix"
    launches = launches_timed
    clock_window = "timed region"
    if t_clk1 - t_clk0 < 0.35:
        clock_window = "timed region + the same load continued untimed to 0.45 s (the region is shorter than the 100 ms sampling period x 4)"
        # the timed region is shorter than a few 100 ms sampling periods: keep the same load running (untimed) until
        # the sampler has seen it, and report those samples (the state is advanced, nothing below depends on it)
        while time.time() - t_clk0 < 0.45:
            sim.tick_n(ts, max(1, args.steps // 4))
            sim.sync()
        t_clk1 = sampler.mark()
    clocks = sampler.stop(t_clk0, t_clk1)
    clocks["window"] = clock_window
    sim.enable_timing(False)
    ms_per_step = ms / args.steps
    value = n * args.steps / (ms * 1e-3)

    ph = dict(phases)
    # steady state on the counting path: the previous tick's force pass has already produced this tick's keys and cell
    # counts (DESIGN.md "speculation"), so no predict+key kernel runs and its phase is empty (faster than 5x the HBM peak
    # could move its 28 B/particle). Its algorithmic bytes then belong to the force pass, which does that work.
    algo = dict(ALGO_BYTES)
    key_fused = sort_path == "counting" and ph.get("predict+key+hist", 1.0) * 1e-3 < 0.2 * ALGO_BYTES["predict+key+hist"] * n / (peak * 1e9)
    if key_fused:
        algo["force+integrate"] += algo.pop("predict+key+hist")
    dom = max((k_ for k_ in ph if k_ in algo), key=ph.get)
    dom_ms = ph[dom]
    achieved = algo[dom] * n / (dom_ms * 1e-3) / 1e9
    kernel_of = {"force+integrate": "k_force_integrate", "density": "k_density", "sort": "k_onesweep_pass",
                 "predict+key+hist": "k_predict_key_hist", "gather+cells": "k_gather_cells"}
    if sort_path == "counting":  # kernels_count.cuh; the phase names and SURVEY's algorithmic bytes stay the same
        kernel_of.update({"sort": "k_cell_slots", "predict+key+hist": "k_predict_key_count", "gather+cells": "k_rank_gather_cells"})
    roofline = {"bound": "hbm", "kernel": kernel_of[dom],
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "algorithmic_bytes_per_particle": algo[dom], "kernel_ms": dom_ms,
                "key_pass": "emitted by the previous tick's force pass (its 28 B/particle are counted with force+integrate)" if key_fused
                            else "own kernel",
                "share_of_step": dom_ms / ms_per_step,
                "whole_step": {"algorithmic_bytes_per_particle": ALGO_BYTES_TOTAL,
                               "achieved": ALGO_BYTES_TOTAL * n / (ms_per_step * 1e-3) / 1e9,
                               "frac": ALGO_BYTES_TOTAL * n / (ms_per_step * 1e-3) / 1e9 / peak},
                "phases_ms": {k: round(v, 5) for k, v in phases},
                "between_ticks_ms": None if between_ticks_ms is None else round(between_ticks_ms, 5),
                "phases_source": "events inside the timed region" if marks_in_region else
                                 "second pass of the same K steps with per-phase events (the timed region runs without them)",
                "phase_frac_of_peak": {k: round(algo[k] * n / (v * 1e-3) / 1e9 / peak, 4) for k, v in phases if v > 0 and k in algo},
                "phase_bytes": {"note": "phase_frac_of_peak charges SURVEY.md 8(d)'s per-particle model (a 3-pass radix sort) to every phase; "
                                        "the counting path moves far fewer bytes in 'sort' and 'gather+cells', so those model fractions "
                                        "exceed 1 without any kernel exceeding the HBM peak: moved_frac_of_peak uses the bytes actually moved",
                                "algorithm": sort_path, "survey_model_bytes_per_particle": {k: algo[k] for k in algo},
                                "moved_bytes_per_particle": MOVED_BYTES[sort_path],
                                "moved_frac_of_peak": {k: round(MOVED_BYTES[sort_path][k] * n / (v * 1e-3) / 1e9 / peak, 4)
                                                       for k, v in phases if v > 0 and k in MOVED_BYTES[sort_path]}}}
    traffic_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(traffic_path):
        with open(traffic_path) as f:
            tj = json.load(f)
        roofline["traffic"] = tj.get(args.workload, {}).get(roofline["kernel"])  # bytes per launch, ncu (profiles/)

    # ---- the same K steps in FLUID_MATH_FAST (hardware rsqrt / rcp + FMA in the pair terms: <= 1e-5 of sum|terms| from
    # the exact mode, inside the north star's tolerance; keys, order, tables, densities unchanged) ----------------------
    also = None
    if not args.no_also:
        sim.math_mode = 1
        sim.tick_n(ts, 3)
        sim.sync()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sim.enable_timing(True)
        with torch.cuda.stream(stream):
            f0.record(stream)
            sim.tick_n(ts, args.steps)
            f1.record(stream)
        sim.sync()
        fast_ms = f0.elapsed_time(f1) / args.steps
        roofline["fast_math"] = {"ms_per_step": fast_ms, "value": n / (fast_ms * 1e-3),
                                 "force_ms": dict(sim.last_timing()).get("force+integrate")}
        sim.enable_timing(False)
        sim.math_mode = 1 if args.math == "fast" else 0
        sim.tick_n(ts, 1)
        sim.sync()

    # ---- end to end through the public API with host buffers -------------------------------------
    e2e = None
    if not args.no_e2e:
        host_in = torch.empty((n, 8), dtype=torch.float32).pin_memory()
        host_outs = [torch.empty((n, 8), dtype=torch.float32).pin_memory() for _ in range(2)]
        out_views = [h.numpy().view(fluid.PARTICLE_DTYPE).reshape(-1) for h in host_outs]
        init = sim.download_particles()
        host_in.numpy().view(np.uint8).reshape(-1)[:] = init.view(np.uint8).reshape(-1)
        del init
        k = max(2, args.e2e_steps)
        # (a) the calls one after the other: upload, tick, download (which synchronises)
        sim.upload_particles(host_in); sim.tick(ts); sim.download_particles(out=out_views[0])
        ks = 2
        t0 = time.perf_counter()
        for j in range(ks):
            sim.upload_particles(host_in)
            sim.tick(ts)
            sim.download_particles(out=out_views[j & 1])  # synchronises
        dt_serial = time.perf_counter() - t0
        # (b) the same three calls per step through the pipelined variants: every step still uploads its full input
        # state from pinned host memory and downloads its full result, but the upload of step k+1 runs on its own copy
        # stream while the result of step k goes back (PCIe is full duplex); results alternate between two host buffers
        sim.upload_particles_async(host_in); sim.tick(ts); sim.download_particles_async(out_views[0]); sim.sync_transfers()
        t0 = time.perf_counter()
        for j in range(k):
            sim.upload_particles_async(host_in)
            sim.tick(ts)
            sim.download_particles_async(out_views[j & 1])
        sim.sync_transfers()
        dt = time.perf_counter() - t0
        e2e = {"value": n * k / dt, "unit": "particle-steps/s", "h2d_bytes_per_step": n * 32 + 60, "d2h_bytes_per_step": n * 32,
               "steps": k, "ms_per_step": 1e3 * dt / k,
               "note": "upload_particles_async (pinned host AoS) + tick + download_particles_async (pinned host AoS) every step, "
                       "one fluid_sim_sync_transfers after the last step; the upload of the next input overlaps the download of the previous result; "
                       "the timed region includes the pipeline's fill and drain",
               "serial": {"value": n * ks / dt_serial, "ms_per_step": 1e3 * dt_serial / ks, "steps": ks,
                          "note": "upload_particles + tick + download_particles, each call complete before the next"}}
        # what the copies alone cost on this box: the same bytes, same pinned buffers, no tick, no transposition — one
        # direction at a time and both at once (the pipelined loop above cannot be faster than the last figure)
        dev_in = torch.empty((n, 8), dtype=torch.float32, device=f"cuda:{local_rank}")
        dev_out = torch.empty((n, 8), dtype=torch.float32, device=f"cuda:{local_rank}")
        s_up, s_dn = torch.cuda.Stream(device=local_rank), torch.cuda.Stream(device=local_rank)

        def copies(up, down, reps=3):
            torch.cuda.synchronize(local_rank)
            t0 = time.perf_counter()
            for j in range(reps):
                if up:
                    with torch.cuda.stream(s_up):
                        dev_in.copy_(host_in, non_blocking=True)
                if down:
                    with torch.cuda.stream(s_dn):
                        host_outs[j & 1].copy_(dev_out, non_blocking=True)
            torch.cuda.synchronize(local_rank)
            return 1e3 * (time.perf_counter() - t0) / reps

        copies(True, True, 1)
        e2e["copy_floor"] = {"h2d_only_ms": copies(True, False), "d2h_only_ms": copies(False, True), "both_at_once_ms": copies(True, True),
                             "note": "cudaMemcpyAsync of the step's input and output bytes between the same pinned buffers and device memory, nothing else"}
        e2e["copy_floor"]["e2e_over_floor"] = e2e["ms_per_step"] / e2e["copy_floor"]["both_at_once_ms"]
        del dev_in, dev_out
        host_out = None
        del host_in, host_outs, out_views, host_out
        # the reference's own per-frame call pattern, for comparison (extra key, NOT the contract's e2e): the particle
        # state stays on the device as it does in the reference; per frame the host uploads the force field
        # (queue.write_buffer, src/renderer.rs:498-502), calls tick with host TickSettings, and receives the frame
        # (fluid_shader.wgsl at 960 x 540, src/renderer.rs:15, as a headless RGBA8 image in host memory).
        from gpu_fluid_simulation_b200 import field as F
        tw, th = s.texture_size
        field_pinned = torch.zeros((th, tw, 2), dtype=torch.float32).pin_memory()
        field_host = field_pinned.numpy()
        inv = F.inverse_projection(s.size)
        kf = max(3, k)
        sim.upload_force_field(field_host); sim.tick(ts); F.render(sim, inv, 960, 540, want_f32=False, want_rgba8=True)
        t0 = time.perf_counter()
        for _ in range(kf):
            sim.upload_force_field(field_host)
            sim.tick(ts)
            F.render(sim, inv, 960, 540, want_f32=False, want_rgba8=True)  # synchronises
        dtf = time.perf_counter() - t0
        e2e["frame_pattern"] = {"value": n * kf / dtf, "unit": "particle-steps/s", "ms_per_step": 1e3 * dtf / kf, "steps": kf,
                                "h2d_bytes_per_step": int(field_host.nbytes) + 60, "d2h_bytes_per_step": 960 * 540 * 4,
                                "note": "state resident (as in the reference); per step: force-field upload + tick + headless "
                                        "960x540 render downloaded to the host"}

    sim.close()
    # ---- the metric's other sizes (BASELINE.json: "at 1M & 64M particles"), a second or two each --------------------
    if not args.no_also and args.workload == "64m-u":
        also = {}
        for name, wl, warm, steps in (("1m-u", "1m-u", 20, 400), ("1m-d", "1m-d", 20, 200), ("1m-d after 300 ticks", "1m-d", 300, 100), ("1m-d after 400 ticks", "1m-d", 400, 100),
                                      ("default (100k, the reference's scene)", "default", 20, 400)):
            s2, ts2, p2 = make_scene(wl)
            with fluid.FluidSimulation.new(local_rank, s2) as sim2:
                if p2 is not None:
                    sim2.upload_particles(p2)
                sim2.tick_n(ts2, warm)
                sim2.sync()
                st2 = torch.cuda.ExternalStream(sim2.stream, device=local_rank)
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(st2):
                    g0.record(st2)
                    sim2.tick_n(ts2, steps)
                    g1.record(st2)
                sim2.sync()
                ms2 = g0.elapsed_time(g1) / steps
                also[name] = {"particles": s2.particle_count, "ms_per_step": round(ms2, 5), "value": s2.particle_count / (ms2 * 1e-3),
                              "steps": steps, "warmup": warm, "sort_path": "counting" if sim2.last_sort_path == 2 else "radix"}

    cpu = None
    if not args.no_cpu_baseline:
        cpu = run_cpu(args.workload, 8, warmup=1, budget_s=12.0)
        port = run_cpu(args.workload, 8, warmup=1, budget_s=8.0, kind="port") if cpu["kind"] != "port" else cpu
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        cpu["port"] = {k: port[k] for k in ("value", "sample")}  # the same step with a fair CPU sort (hand-written oracle)

    line = {"metric": "particle-steps/sec", "value": value, "unit": "particle-steps/s", "n_gpus": 1, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles": n, "math": args.math, "sort": args.sort, "sort_path": sort_path,
                       "l2": "inputs larger than L2 (no flush)" if n * 64 > 126e6 * 2 else "state fits L2; steps run back to back as in the real simulation loop (no flush)",
                       "parallelism": "1 GPU"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "also": also}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()

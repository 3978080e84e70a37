"""bench.py --gpus N (N > 1), one rank per GPU: the scene is cut into bands of grid rows and every rank drives its band
through libfluid_b200's own multi-GPU step (fluid_multi_create_rank / fluid_multi_tick: csrc/multi.inl). torch.distributed
is only the launcher's plumbing here — it hands rank 0's NCCL id to the other ranks, gathers results for the parity check
and takes the max over the ranks' timings; no torch call sits inside a tick."""
import os
import time

import numpy as np
import torch

from . import _lib, multi, scenes
from .simulation import FluidSimulation


def default_row_bounds(grid_h: int, world: int):
    inner = max(grid_h - 2, 1)
    per = -(-inner // world)
    return [0] + [min(1 + r * per, grid_h) for r in range(1, world)] + [grid_h]


def scene_dims(kind, arg):
    """(settings, tick settings, lattice rows) without generating particles."""
    if kind == "uniform":
        s, ts, _ = scenes.uniform_box(arg, lattice_rows=(0, 0))
        return s, ts, arg
    s, ts, _ = scenes.dam_break(arg[0], arg[1], lattice_rows=(0, 0))
    return s, ts, arg[1]


def make_band(kind, arg, bounds, rank):
    """this rank's particles, generated locally (no rank ever holds the whole scene): a contiguous range of the scene's
    array, so the bands concatenated in rank order ARE the scene's array."""
    _, _, rows = scene_dims(kind, arg)
    r0, r1 = scenes.band_lattice_rows(bounds, rank, rows)
    if kind == "uniform":
        return scenes.uniform_box(arg, lattice_rows=(r0, r1))[2]
    return scenes.dam_break(arg[0], arg[1], lattice_rows=(r0, r1))[2]


def grid_h_of(s):
    return int(np.ceil(np.float32(s.size[1]) / np.float32(s.smoothing_radius))) + 2


def new_rank_sim(dist, rank, world, local_rank, s, **opt):
    """rank 0 makes the NCCL id, the launcher's process group carries the 128 bytes to the others."""
    idt = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        idt = torch.tensor(list(multi.unique_id()), dtype=torch.uint8)
    idt = idt.cuda(local_rank)
    dist.broadcast(idt, 0)
    m = multi.MultiFluidSimulation.new_rank(rank, world, bytes(idt.cpu().tolist()), local_rank, s, **opt)
    m.rank = rank
    return m


def _gather_to_rank0(dist, rank, world, local_rank, arr: np.ndarray):
    """variable-length uint8 gather over the launcher's process group (plumbing; outside any timed region)."""
    n = torch.tensor([arr.nbytes], dtype=torch.int64, device=f"cuda:{local_rank}")
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(x.item()) for x in sizes]
    mx = max(max(sizes), 1)
    buf = torch.zeros(mx, dtype=torch.uint8, device=f"cuda:{local_rank}")
    if arr.nbytes:
        buf[: arr.nbytes] = torch.from_numpy(arr.view(np.uint8).reshape(-1)).to(buf.device)
    out = [torch.empty(mx, dtype=torch.uint8, device=buf.device) for _ in range(world)] if rank == 0 else None
    dist.gather(buf, out, dst=0)
    if rank != 0:
        return None
    return [o[:sz].cpu().numpy() for o, sz in zip(out, sizes)]


def parity_vs_single_gpu(dist, rank, world, local_rank, kind, arg, ticks=3, math_mode=0, **opt):
    """K ticks of the scene through the NCCL path, the bands gathered on rank 0 and compared — every field of every
    particle and the whole start_indices table — with a single-GPU run of the same array. Returns (on rank 0) the
    number of differing fields / cells."""
    s, ts, _ = scene_dims(kind, arg)
    bounds = default_row_bounds(grid_h_of(s), world)
    band = make_band(kind, arg, bounds, rank)
    m = new_rank_sim(dist, rank, world, local_rank, s, row_bounds=bounds, **opt)
    try:
        m.set_math_mode(math_mode)
        transport = multi.TRANSPORT_NAMES.get(m.transport, str(m.transport))
        m.upload_band(band)
        m.tick_n(ts, ticks)
        m.sync()
        got, goff = m.download_band()
        table = m.download_start_indices()
        gw, gh = m.grid
    finally:
        m.close()
    parts = _gather_to_rank0(dist, rank, world, local_rank, got)
    lo, hi = bounds[rank] * gw, bounds[rank + 1] * gw
    tables = _gather_to_rank0(dist, rank, world, local_rank, np.ascontiguousarray(table[lo:hi]))
    inits = _gather_to_rank0(dist, rank, world, local_rank, band)
    res = None
    if rank == 0:
        got_all = np.concatenate([p.view(_lib.PARTICLE_DTYPE) for p in parts])
        table_all = np.concatenate([t.view(np.uint32) for t in tables])
        init = np.concatenate([p.view(_lib.PARTICLE_DTYPE) for p in inits])
        with FluidSimulation.new(local_rank, s) as ref:
            ref.math_mode = math_mode
            ref.upload_particles(init)
            ref.tick_n(ts, ticks)
            want = ref.download_particles()
            want_table = ref.download_start_indices()
        diffs = 0
        if got_all.shape != want.shape:
            diffs = -1
        else:
            for f in ("grid", "predicted_position", "density", "velocity", "position"):
                diffs += int((got_all[f] != want[f]).sum())
            diffs += int((table_all != want_table).sum())
        res = {"scene": f"{kind}:{arg}", "particles": int(want.shape[0]), "ticks": ticks, "fields_differing": diffs, "transport": transport}
    flag = torch.tensor([0 if res is None else res["fields_differing"]], dtype=torch.int64, device=f"cuda:{local_rank}")
    dist.broadcast(flag, 0)
    return res, int(flag.item())


def run_multi_bench(args, rank, world, local_rank, workloads, algo_bytes, peak, peak_src, sampler_cls):
    import torch.distributed as dist
    kind, arg = workloads[args.workload]
    if kind == "default":
        raise SystemExit("--gpus > 1 benches the uniform-box and dam-break workloads")
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    math_mode = 1 if args.math == "fast" else 0

    # ---- parity first: the NCCL path against one GPU, on a uniform box and on a dam break -------------------------
    parity = []
    if not args.no_parity:
        for pk, pa, pt, popt in (("uniform", 2000, 3, {}), ("dam", (500, 2000), 3, {"window_near": 65536})):
            res, bad = parity_vs_single_gpu(dist, rank, world, local_rank, pk, pa, ticks=pt, math_mode=math_mode, **popt)
            if bad != 0:
                raise SystemExit(f"multi-GPU parity check failed: {res}")
            parity.append(res)

    s, ts, lattice_rows = scene_dims(kind, arg)
    n_total = s.particle_count
    bounds = default_row_bounds(grid_h_of(s), world)
    band = make_band(kind, arg, bounds, rank)
    per_row = max(1, n_total // max(grid_h_of(s) - 2, 1))
    opt = {"row_bounds": bounds}
    if kind == "dam":
        opt["window_near"] = max(64 * per_row, 65536)   # rows fill up as the column collapses
        opt["capacity"] = min(n_total, int(n_total / world * 3) + 65536)
    m = new_rank_sim(dist, rank, world, local_rank, s, **opt)
    m.set_math_mode(math_mode)
    m.upload_band(band)
    n_start = int(band.shape[0])
    del band
    stream = torch.cuda.ExternalStream(m.stream(0), device=local_rank)
    m.tick_n(ts, args.warmup)
    m.sync()
    # the clock sampler (an nvidia-smi subprocess on rank 0) starts BEFORE the barrier (its start-up must not delay rank 0's
    # first exchange inside the timed region)
    sampler = sampler_cls(local_rank)
    if rank == 0:
        sampler.start()
    dist.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    launches0 = m.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    m.tick_n(ts, args.steps)   # ONE library call: K whole ticks enqueued back to back, nothing read back
    e1.record(stream)
    m.sync()
    torch.cuda.synchronize()
    launches_timed = m.launch_count - launches0
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local_rank}")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    # per-phase times from a second pass of the same K steps with the library's event marks on (8 cudaEventRecord per
    # tick cost ~20 us: the timed region above runs without them)
    m.enable_timing(True)
    m.tick_n(ts, args.steps)
    m.sync()
    phases = m.last_timing(0)
    m.enable_timing(False)
    launches = torch.tensor([launches_timed], dtype=torch.int64, device=f"cuda:{local_rank}")
    dist.all_reduce(launches)
    counts = m.particle_counts()
    assert sum(counts) == n_total, "particles were lost or duplicated in migration"

    # ---- end to end: every rank uploads its band from pinned host memory, ticks, downloads it --------------------
    e2e = None
    if not args.no_e2e:
        k = max(2, args.e2e_steps)
        got, _ = m.download_band()
        n0 = int(got.shape[0])
        cap = n0 + 4 * per_row + 4096          # a tick changes a band's size by the few particles that cross its two borders
        host_in = torch.empty((max(n0, 1), 8), dtype=torch.float32).pin_memory()
        view_in = host_in.numpy().view(_lib.PARTICLE_DTYPE).reshape(-1)
        view_in[:n0] = got
        outs = [torch.empty((cap, 8), dtype=torch.float32).pin_memory() for _ in range(2)]
        views = [o.numpy().view(_lib.PARTICLE_DTYPE).reshape(-1) for o in outs]
        infos = [torch.zeros(2, dtype=torch.int32).pin_memory() for _ in range(2)]
        del got
        # per step, through the library's pipelined band transfers: upload this rank's band (pinned AoS), tick (with its
        # exchange), download the band (pinned AoS); the upload of step k+1 runs under the download of step k
        m.upload_band_async(view_in[:n0]); m.tick(ts); m.download_band_async(views[0], infos[0].numpy().view(np.uint32)); m.sync_transfers()
        dist.barrier()
        t0 = time.perf_counter()
        for j in range(k):
            m.upload_band_async(view_in[:n0])
            m.tick(ts)
            m.download_band_async(views[j & 1], infos[j & 1].numpy().view(np.uint32))
        m.sync_transfers()
        dt_local = time.perf_counter() - t0
        dist.barrier()
        dt = torch.tensor([dt_local], dtype=torch.float64, device=f"cuda:{local_rank}")
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        n_after = int(infos[(k - 1) & 1][0])
        assert 0 < n_after <= cap, "the download window was too small for the band"
        # the pipelined download must be what the plain one returns
        chk, _ = m.download_band()
        assert chk.shape[0] == n_after and np.array_equal(chk, views[(k - 1) & 1][:n_after]), "pipelined band download differs from the synchronous one"
        tot = torch.tensor([n0 * 32, cap * 32 + 8, n_after], dtype=torch.int64, device=f"cuda:{local_rank}")
        dist.all_reduce(tot)
        assert int(tot[2].item()) == n_total, "particles were lost or duplicated in the end-to-end loop"
        e2e = {"value": n_total * k / float(dt.item()), "unit": "particle-steps/s", "h2d_bytes_per_step": int(tot[0].item()),
               "d2h_bytes_per_step": int(tot[1].item()), "steps": k, "ms_per_step": 1e3 * float(dt.item()) / k,
               "rank0_gbs": {"h2d": n0 * 32 * k / dt_local / 1e9, "d2h": cap * 32 * k / dt_local / 1e9},
               "note": "every rank per step: fluid_multi_upload_band_async (its band, pinned host AoS) + fluid_multi_tick (with the exchange) + "
                       "fluid_multi_download_band_async (pinned host AoS); one fluid_multi_sync_transfers after the last step; all ranks "
                       "concurrently, the upload of the next step overlapping the download of the previous one"}
        # what the copies alone cost when every rank moves its band both ways at once (same pinned buffers, no tick, no
        # transposition): the floor of the loop above on this box
        dev_in = torch.empty((max(n0, 1), 8), dtype=torch.float32, device=f"cuda:{local_rank}")
        dev_out = torch.empty((cap, 8), dtype=torch.float32, device=f"cuda:{local_rank}")
        s_up, s_dn = torch.cuda.Stream(device=local_rank), torch.cuda.Stream(device=local_rank)
        floor = {}
        for name, up, down in (("h2d_only_ms", True, False), ("d2h_only_ms", False, True), ("both_at_once_ms", True, True)):
            torch.cuda.synchronize(local_rank)
            dist.barrier()
            t0 = time.perf_counter()
            for j in range(3):
                if up:
                    with torch.cuda.stream(s_up):
                        dev_in.copy_(host_in, non_blocking=True)
                if down:
                    with torch.cuda.stream(s_dn):
                        outs[j & 1].copy_(dev_out, non_blocking=True)
            torch.cuda.synchronize(local_rank)
            t = torch.tensor([(time.perf_counter() - t0) / 3], dtype=torch.float64, device=f"cuda:{local_rank}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            floor[name] = 1e3 * float(t.item())
        floor["note"] = "all ranks at once, max over ranks: cudaMemcpyAsync of each band's bytes between the same pinned buffers and device memory, nothing else"
        floor["e2e_over_floor"] = e2e["ms_per_step"] / floor["both_at_once_ms"]
        e2e["copy_floor"] = floor
        del host_in, outs, views, view_in, dev_in, dev_out

    ms_per_step = ms / args.steps
    ph = dict(phases)
    dom_ms = ph.get("force+integrate", 0.0)
    n_rank = int(counts[rank])
    # as on one GPU: in steady state the force pass also runs the next tick's key pass (no separate key kernel runs: its
    # phase is a few microseconds of header writing), so that pass's 28 B per particle are counted with it
    spec = os.environ.get("FLUID_NO_SPECULATION", "0") != "1" and args.steps > 1
    dom_bytes = algo_bytes["force+integrate"] + (algo_bytes["predict+key+hist"] if spec else 0)
    achieved = dom_bytes * n_rank / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    line = {"metric": "particle-steps/sec", "value": n_total * args.steps / (ms * 1e-3), "unit": "particle-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles": n_total, "math": args.math,
                       "parallelism": f"{world} bands of grid rows, one rank per GPU; the whole step incl. its one exchange per tick "
                                      "(2 ghost rows + migrants per neighbour) runs inside libfluid_b200 (fluid_multi_tick); no host read per tick",
                       "transport": multi.TRANSPORT_NAMES.get(m.transport, str(m.transport)),
                       "row_bounds": bounds, "particles_per_rank_start": None, "particles_per_rank_end": counts,
                       "imbalance_end": round(max(counts) * world / n_total, 4),
                       "l2": "inputs larger than L2 (no flush)" if n_total * 64 / world > 252e6 else "per-rank state fits L2; steps run back to back (no flush)"},
            "parity_vs_single_gpu": {"world": world, "fields_differing": 0 if parity else None, "checks": parity},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches.item()),
            "roofline": {"bound": "hbm", "kernel": "k_force_integrate (rank 0)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_particle": dom_bytes, "kernel_ms": dom_ms,
                         "key_pass": "emitted by the previous tick's force pass (its 28 B/particle are counted with force+integrate)" if spec else "own kernel",
                         "share_of_step": dom_ms / ms_per_step if ms_per_step else None,
                         "phases_ms_rank0": {k_: round(v, 5) for k_, v in phases},
                         "phases_source": "second pass of the same K steps with per-phase events (the timed region runs without them)"},
            "cpu_baseline": None}
    line["config"]["particles_per_rank_start"] = [int(x) for x in _all_gather_int(dist, n_start, world, local_rank)]
    m.close()
    dist.barrier()
    return line


def _all_gather_int(dist, v, world, local_rank):
    t = torch.tensor([v], dtype=torch.int64, device=f"cuda:{local_rank}")
    out = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    return [int(x.item()) for x in out]

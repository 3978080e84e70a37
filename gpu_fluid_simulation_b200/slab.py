"""Slab (multi-GPU) mode: the domain is cut into bands of grid rows, one band per rank.

One process per GPU; the device passes are libfluid_b200's fluid_slab_* phases (csrc/slab.inl); the three exchanges
per tick move through torch.distributed (NCCL over NVLink) on zero-copy views of the handle's device arrays:

  1. migration   particles whose predicted position left the band -> their new owner (all_to_all: counts, payload).
                 Immigrants from lower ranks are placed before the natives, from higher ranks after them, in source
                 order, so the stable sort reproduces the single-GPU tie order exactly.
  2. halo        first / last grid row of (pred, vel) -> lower / upper neighbour (P2P send/recv)
  3. halo rho    density of the same rows, after the density pass

Keys stay global, so the concatenation of the ranks' sorted arrays IS the single-GPU sorted array, and every rank
computes its particles with the neighbours in the same order as one GPU would: results are bit-identical to a
single-GPU tick (tests/test_gpu_slab.py).

`SlabRank` holds the per-rank phases and is transport-agnostic; `LocalDriver` runs several ranks in one process on
one device (tests, no NCCL), `DistDriver` runs one rank per process over torch.distributed.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .simulation import FluidSimulation, SimulationSettings, TickSettings

NO_QUIRK = 0xFFFFFFFF


def row_bounds(grid_h: int, world: int):
    """bands of grid rows: rows 1..grid_h-2 can hold particles; the empty border rows go to the edge ranks."""
    inner = max(grid_h - 2, 1)
    per = -(-inner // world)
    b = [0]
    for r in range(1, world):
        b.append(min(1 + r * per, grid_h))
    b.append(grid_h)
    return b


class _DeviceArray:
    """__cuda_array_interface__ wrapper so torch can view library-owned device memory without a copy."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def _view(ptr, shape, typestr, device):
    return torch.as_tensor(_DeviceArray(ptr, shape, typestr), device=device)


class SlabRank:
    """one rank's band. All methods only enqueue work on the handle's stream unless they say they synchronise."""

    def __init__(self, device, settings: SimulationSettings, rank: int, world: int, capacity: int, pad: int,
                 emig_cap: int = 1 << 16, math_mode: int = 0):
        self.rank, self.world = rank, world
        self.settings = settings
        self.sim = FluidSimulation.new(device, SimulationSettings(particle_count=2, particle_spacing=settings.particle_spacing,
                                                                  smoothing_radius=settings.smoothing_radius, size=settings.size,
                                                                  texture_size=settings.texture_size))
        self.sim.math_mode = math_mode
        self.lib = self.sim._lib
        self.h = self.sim._h
        self.grid_w, self.grid_h = self.sim.grid_w, self.sim.grid_h
        self.bounds = row_bounds(self.grid_h, world)
        rb = (C.c_uint32 * (world + 1))(*self.bounds)
        _lib.check(self.lib.fluid_slab_configure(self.h, rank, world, rb, int(capacity), int(pad), int(emig_cap)))
        v = _lib.SlabView()
        _lib.check(self.lib.fluid_slab_view(self.h, C.byref(v)))
        self.pad, self.capacity, self.emig_cap = v.pad, v.capacity, v.emig_cap
        self.device = torch.device("cuda", device)
        total = self.capacity + 2 * self.pad
        dv = self.device
        self.pos_state = _view(v.pos_state, (total, 2), "<f4", dv)
        self.vel_state = _view(v.vel_state, (total, 2), "<f4", dv)
        self.pred_sorted = _view(v.pred_sorted, (total, 2), "<f4", dv)
        self.vel_sorted = _view(v.vel_sorted, (total, 2), "<f4", dv)
        self.rho = _view(v.rho, (total,), "<f4", dv)
        self.emig_list = _view(v.emig_list, (self.emig_cap,), "<i8", dv)
        self.meta = _view(v.meta, (64,), "<i4", dv)  # words 0..4 band metadata, word 8 emigrant counter, 16.. per destination
        self.dest_counts = self.meta[16:16 + world]
        self.dyn = _view(v.dyn, (8,), "<i4", dv)  # device-resolved halo counts / offsets / error bits (SlabDyn)
        self.stream = torch.cuda.ExternalStream(self.sim.stream, device=dv)
        self.n = 0
        self.n_emig = 0
        self.n_below = self.n_above = 0
        self.halo_below = self.halo_above = 0

    def stream_ctx(self):
        """torch ops and NCCL calls issued inside are ordered on the handle's CUDA stream."""
        return torch.cuda.stream(self.stream)

    # ---- setup -------------------------------------------------------------------------------------------
    def owns_rows(self):
        return self.bounds[self.rank], self.bounds[self.rank + 1]

    def upload_band(self, particles: np.ndarray):
        """particles of this band in global array order (AoS)."""
        self.n = int(particles.shape[0])
        _lib.check(self.lib.fluid_sim_set_particle_count(self.h, self.n))
        if self.n:
            self.sim.upload_particles(np.ascontiguousarray(particles))
        self.sim.sync()

    def download_band(self) -> np.ndarray:
        out = np.empty(self.n, dtype=_lib.PARTICLE_DTYPE)
        if self.n:
            _lib.check(self.lib.fluid_sim_download_particles(self.h, _lib.ptr(out)))
        return out

    # ---- phase 1: keys of the natives, emigrant list --------------------------------------------------------
    def begin_tick(self, ts: TickSettings):
        c = ts.to_c()
        _lib.check(self.lib.fluid_slab_begin_tick(self.h, C.byref(c)))

    def send_counts(self) -> torch.Tensor:
        """int64[world] on the device: emigrants per destination rank, counted by the key pass (no host sync)."""
        with torch.cuda.stream(self.stream):
            return self.dest_counts.to(torch.int64)

    def pack(self, send_counts_host):
        """payload float32[m, 4] = (pos, vel) of the emigrants grouped by destination, source order inside a group
        (the library sorts the emigrant list and gathers the payload)."""
        m = int(sum(send_counts_host))
        self.n_emig = m
        if m > self.emig_cap:
            raise _lib.FluidError(f"rank {self.rank}: {m} emigrants exceed emig_cap {self.emig_cap}")
        with torch.cuda.stream(self.stream):
            payload = torch.empty((m, 4), dtype=torch.float32, device=self.device)
            if m:
                _lib.check(self.lib.fluid_slab_pack(self.h, m, _lib.ptr(payload)))
            return payload

    # ---- phase 2: immigrants in -------------------------------------------------------------------------------
    def unpack(self, recv: torch.Tensor, recv_counts_host):
        """recv float32[k, 4] ordered by source rank."""
        nb = int(sum(recv_counts_host[: self.rank]))
        na = int(sum(recv_counts_host[self.rank + 1:]))
        if nb > self.pad or na > self.pad:
            raise _lib.FluidError(f"rank {self.rank}: {nb}+{na} immigrants exceed the pad {self.pad}")
        self.n_below, self.n_above = nb, na
        if nb + na:
            recv = recv.contiguous()
            _lib.check(self.lib.fluid_slab_unpack(self.h, _lib.ptr(recv), nb, na))
            self._recv_keepalive = recv  # the copy kernel is only enqueued

    # ---- phases 3+4: sort, gather, metadata ---------------------------------------------------------------------
    def sort(self):
        _lib.check(self.lib.fluid_slab_sort(self.h, self.n_below, self.n_above, self.n_emig))
        self.n = self.n + self.n_below + self.n_above - self.n_emig

    def meta_tensor(self) -> torch.Tensor:
        """int64[5] on the device: n_owned, first-row count, last-row count, key of first owned, its table entry."""
        with torch.cuda.stream(self.stream):
            return self.meta[:5].to(torch.int64) & 0xFFFFFFFF

    # ---- phase 5: halo rows ------------------------------------------------------------------------------------------
    def halo_out(self, first_count, last_count):
        """(to_below, to_above): lists of contiguous views [pred, vel] of the band's first / last grid row."""
        P, n = self.pad, self.n
        self._first, self._last = int(first_count), int(last_count)
        lo = slice(P, P + self._first)
        hi = slice(P + n - self._last, P + n)
        return [self.pred_sorted[lo], self.vel_sorted[lo]], [self.pred_sorted[hi], self.vel_sorted[hi]]

    def halo_in(self, below_count, above_count):
        P, n = self.pad, self.n
        if below_count > self.pad or above_count > self.pad:
            raise _lib.FluidError(f"rank {self.rank}: halo rows ({below_count}, {above_count}) exceed the pad {self.pad}")
        self.halo_below, self.halo_above = int(below_count), int(above_count)
        lo = slice(P - self.halo_below, P)
        hi = slice(P + n, P + n + self.halo_above)
        return [self.pred_sorted[lo], self.vel_sorted[lo]], [self.pred_sorted[hi], self.vel_sorted[hi]]

    def rho_out(self):
        P, n = self.pad, self.n
        return self.rho[P:P + self._first], self.rho[P + n - self._last:P + n]

    def rho_in(self):
        P, n = self.pad, self.n
        return self.rho[P - self.halo_below:P], self.rho[P + n:P + n + self.halo_above]

    # ---- device-resolved protocol: no host read between the sort and the force pass ---------------------------------
    def meta_raw(self) -> torch.Tensor:
        """int32[8] view of this rank's band metadata on the device (n, first-row count, last-row count, first key, entry)."""
        return self.meta[:8]

    def halo_windows(self):
        """fixed pad-sized windows for the halo exchange of (pred, vel): (to_below, to_above, from_below, from_above),
        each a list [pred, vel]. My first `pad` owned particles go below (their first row is my first row), my last
        `pad` go above; the neighbour's window lands right-aligned before / directly after my owned range. Only the
        halo-row part of a window is ever read (counts are resolved on the device)."""
        P, n, pad = self.pad, self.n, self.pad
        out_lo, out_hi = slice(P, P + pad), slice(P + n - pad, P + n)
        in_lo, in_hi = slice(P - pad, P), slice(P + n, P + n + pad)
        return ([self.pred_sorted[out_lo], self.vel_sorted[out_lo]], [self.pred_sorted[out_hi], self.vel_sorted[out_hi]],
                [self.pred_sorted[in_lo], self.vel_sorted[in_lo]], [self.pred_sorted[in_hi], self.vel_sorted[in_hi]])

    def rho_windows(self):
        P, n, pad = self.pad, self.n, self.pad
        return (self.rho[P:P + pad], self.rho[P + n - pad:P + n], self.rho[P - pad:P], self.rho[P + n:P + n + pad])

    def cells_density_dev(self, allmeta: torch.Tensor):
        """allmeta: int32[world * 8] on the device (all_gather of meta_raw)."""
        self._allmeta_keepalive = allmeta
        _lib.check(self.lib.fluid_slab_cells_density_dev(self.h, _lib.ptr(allmeta)))

    def error_row(self) -> torch.Tensor:
        """int64[world] on the device: element 0 = latched error bits of the device-resolved protocol."""
        with torch.cuda.stream(self.stream):
            row = torch.zeros(self.world, dtype=torch.int64, device=self.device)
            row[0:1] = self.dyn[7:8].to(torch.int64)
            return row

    def raise_on_error_bits(self, bits: int):
        if bits & 1:
            raise _lib.FluidError(f"rank {self.rank}: a halo row is larger than the pad {self.pad} (previous tick)")

    # ---- phases 6+7 ---------------------------------------------------------------------------------------------------
    def cells_density(self, global_offset, quirk_c0, quirk_stale_global):
        """quirk_*: lowest occupied cell of the whole domain and its stale start_indices entry (a GLOBAL sorted index)."""
        if quirk_c0 == NO_QUIRK:
            stale_ext = 0
        else:
            ext_begin = self.pad - self.halo_below
            s = int(quirk_stale_global) - int(global_offset) + self.pad
            stale_ext = min(max(s, ext_begin), 0xFFFFFFFF)
        _lib.check(self.lib.fluid_slab_cells_density(self.h, self.halo_below, self.halo_above, int(global_offset),
                                                     int(quirk_c0) & 0xFFFFFFFF, int(stale_ext)))

    def force(self):
        _lib.check(self.lib.fluid_slab_force(self.h))

    def sync(self):
        self.sim.sync()

    def close(self):
        self.sim.close()


def _quirk_from_metas(metas):
    """metas: int array [world, 5]. The lowest occupied cell belongs to the first non-empty rank."""
    for r in range(metas.shape[0]):
        if metas[r, 0] > 0:
            return int(metas[r, 3]), int(metas[r, 4])
    return NO_QUIRK, 0


class LocalDriver:
    """several SlabRanks in ONE process on one device, exchanging through device copies (tests; no NCCL).
    Ranks run one after another, so nothing ever waits on another rank's kernel."""

    def __init__(self, ranks):
        self.ranks = ranks
        self.world = len(ranks)

    def _sync(self):
        for r in self.ranks:
            r.sync()
        torch.cuda.synchronize()

    def tick(self, ts: TickSettings):
        R, W = self.ranks, self.world
        for r in R:
            r.begin_tick(ts)
        counts = torch.stack([r.send_counts() for r in R]).cpu().numpy()  # [src, dst]; synchronises
        payloads = [r.pack(counts[i]) for i, r in enumerate(R)]
        self._sync()
        for d, r in enumerate(R):
            parts = []
            for s_ in range(W):
                off = int(counts[s_, :d].sum())
                parts.append(payloads[s_][off:off + int(counts[s_, d])])
            r.unpack(torch.cat(parts) if parts else payloads[0][:0], counts[:, d])
        self._sync()
        for r in R:
            r.sort()
        metas = torch.stack([r.meta_tensor() for r in R]).cpu().numpy()
        prefix = np.concatenate([[0], np.cumsum(metas[:, 0])])
        c0, stale = _quirk_from_metas(metas)
        outs = [r.halo_out(metas[i, 1], metas[i, 2]) for i, r in enumerate(R)]
        ins = [r.halo_in(metas[i - 1, 2] if i > 0 else 0, metas[i + 1, 1] if i + 1 < W else 0) for i, r in enumerate(R)]
        self._sync()
        for i in range(W):
            if i > 0:
                for dst, src in zip(ins[i][0], outs[i - 1][1]):
                    dst.copy_(src)
            if i + 1 < W:
                for dst, src in zip(ins[i][1], outs[i + 1][0]):
                    dst.copy_(src)
        self._sync()
        for i, r in enumerate(R):
            r.cells_density(int(prefix[i]), c0, stale)
        self._sync()
        rho_o = [r.rho_out() for r in R]
        rho_i = [r.rho_in() for r in R]
        for i in range(W):
            if i > 0:
                rho_i[i][0].copy_(rho_o[i - 1][1])
            if i + 1 < W:
                rho_i[i][1].copy_(rho_o[i + 1][0])
        self._sync()
        for r in R:
            r.force()
        self._sync()
        self.last_metas = metas


class DistDriver:
    """one SlabRank per process; exchanges over torch.distributed (NCCL on GPUs, gloo in CPU tests of the protocol)."""

    def __init__(self, rank_obj, group=None, device_resolved=None):
        import torch.distributed as dist
        self.dist = dist
        self.r = rank_obj
        self.group = group
        # device-resolved protocol (one host read per tick instead of two) when the rank object supports it
        self.device_resolved = hasattr(rank_obj, "cells_density_dev") if device_resolved is None else bool(device_resolved)
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.exchange_bytes = 0
        self.profile = None  # dict name -> seconds when section profiling is on (scripts/slab_profile.py); serialises the step

    def _section(self, name):
        if self.profile is None:
            return
        import time
        torch.cuda.synchronize()
        now = time.perf_counter()
        self.profile[name] = self.profile.get(name, 0.0) + now - self._t_last
        self._t_last = now

    def tick(self, ts: TickSettings):
        dist, r, W, me = self.dist, self.r, self.world, self.rank
        if self.profile is not None:
            import time
            torch.cuda.synchronize()
            self._t_last = time.perf_counter()
        with r.stream_ctx():
            r.begin_tick(ts)
            self._section("begin_tick")
            # ---- migration: counts, then payload ----
            sc = r.send_counts()
            rc = torch.empty_like(sc)
            self._section("send_counts")
            dist.all_to_all_single(rc, sc, group=self.group)
            rows = [sc, rc] + ([r.error_row()] if self.device_resolved else [])
            both = torch.stack(rows).cpu().numpy()  # synchronises: the only host read of a device-resolved tick
            send_counts, recv_counts = both[0], both[1]
            if self.device_resolved:
                r.raise_on_error_bits(int(both[2][0]))
            self._section("a2a_counts+host")
            payload = r.pack(send_counts)
            self._section("pack")
            recv = torch.empty((int(recv_counts.sum()), 4), dtype=torch.float32, device=r.device)
            if int(send_counts.sum()) or int(recv_counts.sum()) or True:
                dist.all_to_all_single(recv, payload, output_split_sizes=[int(x) for x in recv_counts],
                                       input_split_sizes=[int(x) for x in send_counts], group=self.group)
            self._section("a2a_payload")
            r.unpack(recv, recv_counts)
            self._section("unpack")
            r.sort()
            self._section("sort")
            if self.device_resolved:
                self._tick_tail_device_resolved(send_counts, recv_counts)
                return
            # ---- band metadata of every rank ----
            m = r.meta_tensor()
            allm = torch.empty(W * m.numel(), dtype=m.dtype, device=r.device)
            dist.all_gather_into_tensor(allm, m, group=self.group)
            metas = allm.view(W, -1).cpu().numpy()  # synchronises
            self._section("metas")
            prefix = np.concatenate([[0], np.cumsum(metas[:, 0])])
            c0, stale = _quirk_from_metas(metas)
            # ---- halo rows of (pred, vel) ----
            to_below, to_above = r.halo_out(metas[me, 1], metas[me, 2])
            from_below, from_above = r.halo_in(metas[me - 1, 2] if me > 0 else 0, metas[me + 1, 1] if me + 1 < W else 0)
            ops = []
            for k in range(2):
                if me > 0 and to_below[k].shape[0]:
                    ops.append(dist.P2POp(dist.isend, to_below[k], me - 1, group=self.group))
                if me + 1 < W and to_above[k].shape[0]:
                    ops.append(dist.P2POp(dist.isend, to_above[k], me + 1, group=self.group))
                if me > 0 and from_below[k].shape[0]:
                    ops.append(dist.P2POp(dist.irecv, from_below[k], me - 1, group=self.group))
                if me + 1 < W and from_above[k].shape[0]:
                    ops.append(dist.P2POp(dist.irecv, from_above[k], me + 1, group=self.group))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            self._section("halo_exchange")
            r.cells_density(int(prefix[me]), c0, stale)
            self._section("cells_density")
            # ---- halo rho ----
            ro_b, ro_a = r.rho_out()
            ri_b, ri_a = r.rho_in()
            ops = []
            if me > 0 and ro_b.shape[0]:
                ops.append(dist.P2POp(dist.isend, ro_b, me - 1, group=self.group))
            if me + 1 < W and ro_a.shape[0]:
                ops.append(dist.P2POp(dist.isend, ro_a, me + 1, group=self.group))
            if me > 0 and ri_b.shape[0]:
                ops.append(dist.P2POp(dist.irecv, ri_b, me - 1, group=self.group))
            if me + 1 < W and ri_a.shape[0]:
                ops.append(dist.P2POp(dist.irecv, ri_a, me + 1, group=self.group))
            if ops:
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
            self._section("rho_exchange")
            r.force()
            self._section("force")
        self.last_metas = metas
        self.exchange_bytes = int(16 * (send_counts.sum() + recv_counts.sum()) + 20 * (r.halo_below + r.halo_above + r._first + r._last))


def _p2p(dist, group, me, W, to_below, to_above, from_below, from_above):
    """one batched exchange with both neighbours; every argument is a list of tensors (or a tensor)."""
    as_list = lambda x: list(x) if isinstance(x, (list, tuple)) else [x]  # noqa: E731
    ops = []
    for k in range(len(as_list(to_below))):
        if me > 0:
            ops.append(dist.P2POp(dist.isend, as_list(to_below)[k], me - 1, group=group))
        if me + 1 < W:
            ops.append(dist.P2POp(dist.isend, as_list(to_above)[k], me + 1, group=group))
        if me > 0:
            ops.append(dist.P2POp(dist.irecv, as_list(from_below)[k], me - 1, group=group))
        if me + 1 < W:
            ops.append(dist.P2POp(dist.irecv, as_list(from_above)[k], me + 1, group=group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()


def _tick_tail_device_resolved(self, send_counts, recv_counts):
    """after the sort: metadata all_gather (stays on the device), fixed-size halo windows, cells + density, rho windows,
    force. Nothing here reads from the device."""
    dist, r, W, me = self.dist, self.r, self.world, self.rank
    m = r.meta_raw()
    allm = torch.empty(W * m.numel(), dtype=m.dtype, device=r.device)
    dist.all_gather_into_tensor(allm, m, group=self.group)
    self._section("metas")
    to_below, to_above, from_below, from_above = r.halo_windows()
    _p2p(dist, self.group, me, W, to_below, to_above, from_below, from_above)
    self._section("halo_exchange")
    r.cells_density_dev(allm)
    self._section("cells_density")
    ro_b, ro_a, ri_b, ri_a = r.rho_windows()
    _p2p(dist, self.group, me, W, ro_b, ro_a, ri_b, ri_a)
    self._section("rho_exchange")
    r.force()
    self._section("force")
    self.last_metas = None
    pad = r.pad
    self.exchange_bytes = int(16 * (send_counts.sum() + recv_counts.sum()) + 20 * pad * ((me > 0) + (me + 1 < W)) * 2)


DistDriver._tick_tail_device_resolved = _tick_tail_device_resolved


def split_particles(particles: np.ndarray, settings: SimulationSettings, bounds):
    """initial distribution: by the grid row of the CURRENT predicted position's key, preserving array order.
    (Ownership is re-derived from the keys of each tick, so any consistent initial split works; this one makes the
    first tick migration-free for particles at rest.)"""
    h = np.float32(settings.smoothing_radius)
    half = np.float32(settings.size[1]) * np.float32(0.5)
    row = np.floor((particles["predicted_position"][:, 1] + half) / h).astype(np.int64) + 1
    owner = np.searchsorted(np.asarray(bounds[1:-1]), row, side="right")
    return [particles[owner == r] for r in range(len(bounds) - 1)]


# ------------------------------------------------------------------------------------------------------------------
# bench.py --gpus N (N > 1): the uniform-box workload partitioned into y-slabs, one rank per GPU, NCCL exchanges
# ------------------------------------------------------------------------------------------------------------------
def band_of_uniform_box(n_side, rank, world, spacing=0.1, h=0.2):
    """this rank's particles of the uniform-box scene, generated locally (no rank ever holds the whole scene).
    Lattice rows 2k, 2k+1 (spacing = h/2, jitter < spacing/2) lie in grid row k+1."""
    from . import scenes
    settings, ts, _ = scenes.uniform_box(n_side, spacing, h, lattice_rows=(0, 0))
    gh = int(np.ceil(np.float32(settings.size[1]) / np.float32(h))) + 2
    b = row_bounds(gh, world)
    lo, hi = max(b[rank], 1), b[rank + 1]
    r0, r1 = min(2 * (lo - 1), n_side), min(2 * (hi - 1), n_side)
    _, _, p = scenes.uniform_box(n_side, spacing, h, lattice_rows=(r0, max(r0, r1)))
    return settings, ts, p


def run_slab_bench(args, rank, world, local_rank, workloads, algo_bytes, peak, peak_src, sampler_cls):
    import time
    import torch.distributed as dist
    kind, arg = workloads[args.workload]
    if kind != "uniform":
        raise SystemExit("--gpus > 1 benches the uniform-box workloads (BASELINE.json config 5)")
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    settings, ts, band = band_of_uniform_box(arg, rank, world)
    n_total = settings.particle_count
    per_row = 4 * (int(np.ceil(settings.size[0] / settings.smoothing_radius)) + 2)  # ~4 particles per cell
    pad = max(4 * per_row, 1 << 14)
    cap = int(band.shape[0] * 1.1) + 8 * per_row
    r = SlabRank(local_rank, settings, rank, world, capacity=cap, pad=pad, emig_cap=max(1 << 16, 8 * per_row),
                 math_mode=1 if args.math == "fast" else 0)
    r.upload_band(band)
    del band
    drv = DistDriver(r)
    for _ in range(args.warmup):
        drv.tick(ts)
    r.sync()
    # the clock sampler (an nvidia-smi subprocess on rank 0) starts BEFORE the barrier: started after it, its start-up
    # (tens of ms on some hosts) delayed rank 0's first collective while the other ranks' start events were already
    # recorded, and the max over ranks charged that skew to the timed region (3 ms per step at 20 steps)
    sampler = sampler_cls(local_rank)
    if rank == 0:
        sampler.start()
    r.sim.enable_timing(True)
    drv.tick(ts)  # one more untimed tick with the event marks on (creates the events of the first ring slot)
    r.sync()
    dist.barrier()
    torch.cuda.synchronize()
    dist.barrier()
    launches0 = r.sim.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(r.stream)
    for _ in range(args.steps):
        drv.tick(ts)
    e1.record(r.stream)
    r.sync()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=r.device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    clocks = sampler.stop() if rank == 0 else None
    phases = r.sim.last_timing()
    r.sim.enable_timing(False)
    launches = torch.tensor([r.sim.launch_count - launches0], dtype=torch.int64, device=r.device)
    owned = torch.tensor([r.n], dtype=torch.int64, device=r.device)
    dist.all_reduce(launches)
    all_owned = [torch.zeros_like(owned) for _ in range(world)]
    dist.all_gather(all_owned, owned)
    assert int(sum(int(x.item()) for x in all_owned)) == n_total, "particles were lost or duplicated in migration"
    exch = torch.tensor([drv.exchange_bytes], dtype=torch.int64, device=r.device)
    dist.all_reduce(exch)

    # ---- end to end: every rank uploads its band from pinned host memory, ticks, downloads it ----
    e2e = None
    if not args.no_e2e:
        k = max(1, args.e2e_steps)
        host = torch.empty((r.capacity, 8), dtype=torch.float32).pin_memory()
        view = host.numpy().view(_lib.PARTICLE_DTYPE).reshape(-1)
        h2d = d2h = 0
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(k):
            n_now = r.n
            _lib.check(r.lib.fluid_sim_download_particles(r.h, _lib.ptr(host)))  # synchronises
            d2h += n_now * 32
            _lib.check(r.lib.fluid_sim_upload_particles(r.h, _lib.ptr(host)))
            h2d += n_now * 32
            drv.tick(ts)
        r.sync()
        dist.barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=r.device)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        tot = torch.tensor([h2d, d2h], dtype=torch.int64, device=r.device)
        dist.all_reduce(tot)
        e2e = {"value": n_total * k / float(dt.item()), "unit": "particle-steps/s", "h2d_bytes_per_step": int(tot[0].item()) // k,
               "d2h_bytes_per_step": int(tot[1].item()) // k, "steps": k, "ms_per_step": 1e3 * float(dt.item()) / k,
               "note": "every rank: download band to pinned host AoS + upload it back + tick (all exchanges), per step"}
        del host, view

    ms_per_step = ms / args.steps
    ph = dict(phases)
    dom_ms = ph.get("force+integrate", 0.0)
    n_rank = int(all_owned[rank].item())
    achieved = algo_bytes["force+integrate"] * n_rank / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    line = {"metric": "particle-steps/sec", "value": n_total * args.steps / (ms * 1e-3), "unit": "particle-steps/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "particles": n_total, "math": args.math,
                       "parallelism": f"{world} y-slabs of grid rows, one rank per GPU; NCCL all_to_all migration + P2P halo rows",
                       "particles_per_rank": [int(x.item()) for x in all_owned],
                       "l2": "inputs larger than L2 (no flush)" if n_total * 64 / world > 252e6 else "per-rank state fits L2; steps run back to back (no flush)",
                       "exchange_bytes_per_step_all_ranks": int(exch.item())},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches.item()),
            "roofline": {"bound": "hbm", "kernel": "k_force_integrate (rank 0)", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes_per_particle": algo_bytes["force+integrate"], "kernel_ms": dom_ms,
                         "share_of_step": dom_ms / ms_per_step if ms_per_step else None,
                         "phases_ms_rank0_incl_exchange_waits": {k_: round(v, 5) for k_, v in phases}},
            "cpu_baseline": None}
    r.close()
    dist.barrier()
    return line

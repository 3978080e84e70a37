"""CPU twin of the multi-GPU step's protocol (gpu_fluid_simulation_b200/csrc/kernels_band.cuh + multi.inl), built on the
oracle (TEST INFRASTRUCTURE). One BandRankModel per process; the exchange goes through torch.distributed (gloo in the
CPU tests), so the N > 1 path — who sends which particle to whom, in which message layout, how every rank derives the
other ranks' particle counts and its own global offset without asking anybody, how the in-cell order is restored from
the origins, which rows need ghosts — runs with world_size > 1 on a machine without a GPU. The SPH arithmetic of a
rank is the oracle's, applied to its [ghost rows | owned | ghost rows] array.

The layout mirrors the device code on purpose:
  message = header u32[16] {records, error bits, natives, -, E[me][0..8)} + (pos, vel) f32[cap][4] + origin u32[cap]
  BandDyn = own range, goff / goff_prev, n_all[], latched error bits
"""
import numpy as np
import torch

from oracle import oracle as O

GHOST_ROWS = 2
HDR_WORDS = 16
SENTINEL = 0xFFFFFFFF
ERR_WINDOW, ERR_CAPACITY, ERR_COUNT = 1, 2, 4


def default_bounds(grid_h, world):
    inner = max(grid_h - 2, 1)
    per = -(-inner // world)
    return [0] + [min(1 + r * per, grid_h) for r in range(1, world)] + [grid_h]


class BandRankModel:
    def __init__(self, settings, rank, world, bounds=None, window_near=4096, window_far=256, capacity=None):
        self.s, self.rank, self.world = settings, rank, world
        self.gw, self.gh = O.grid_dims(settings)
        self.bounds = list(bounds) if bounds is not None else default_bounds(self.gh, world)
        self.cap = [0 if r == rank else (window_near if abs(r - rank) == 1 else window_far) for r in range(world)]
        self.capacity = capacity
        self.table = np.zeros(self.gw * self.gh, dtype=np.uint32)  # start_indices, GLOBAL sorted indices
        tw, th = settings.texture_size
        self.texture = np.zeros((th, tw, 2), dtype=np.float32)
        self.tick_count = 0
        self.error = 0
        self.pos = np.zeros((0, 2), np.float32)
        self.vel = np.zeros((0, 2), np.float32)
        self.origin_explicit = None
        self.goff = 0
        self.n_all = np.zeros(world, np.int64)
        self.sorted = np.zeros(0, dtype=O.PARTICLE_DTYPE)
        self.own = (0, 0)

    # ---- rows ----
    def in_ext(self, r, row):
        lo, hi = self.bounds[r], self.bounds[r + 1]
        return (hi > lo) & (row + GHOST_ROWS >= lo) & (row < hi + GHOST_ROWS)

    def _keys(self, pos, vel):
        q = np.zeros(pos.shape[0], dtype=O.PARTICLE_DTYPE)
        q["position"], q["velocity"] = pos, vel
        u = O.make_uniform(self.s, self.ts, self.tick_count)
        u.particle_count = q.shape[0]
        O.predict_next_position(u, q)
        O.create_spatial_lookup(u, q)
        return q

    # ---- upload: any subset of the scene, with the particles' indices in the global array ----
    def upload(self, particles, origins, n_all):
        self.pos = np.array(particles["position"], np.float32)
        self.vel = np.array(particles["velocity"], np.float32)
        self.origin_explicit = np.array(origins, np.uint32)
        self.n_all = np.array(n_all, np.int64)
        self.goff = int(self.n_all[: self.rank].sum())
        self.own = (0, self.pos.shape[0])

    # ---- 1. key pass over the natives: outboxes + this rank's row of the migration matrix ----
    def key_pass(self, ts):
        self.ts = ts
        self.tick_count += 1
        n_nat = self.pos.shape[0]
        q = self._keys(self.pos, self.vel)
        row = (np.minimum(q["grid"], self.gw * self.gh - 1) // self.gw).astype(np.int64)
        origin = self.origin_explicit if self.origin_explicit is not None else (self.goff + np.arange(n_nat)).astype(np.uint32)
        self._nat_keys = np.where(self.in_ext(self.rank, row), q["grid"], SENTINEL).astype(np.uint32)
        self._nat_origin = origin
        msgs, e_row = [], np.zeros(8, np.uint32)
        for d in range(self.world):
            if d == self.rank:
                msgs.append(None)
                continue
            sel = np.nonzero(self.in_ext(d, row))[0]
            e_row[d] = int(((row[sel] >= self.bounds[d]) & (row[sel] < self.bounds[d + 1])).sum())
            msgs.append(sel)
        self.e_sent = e_row
        out = []
        for d in range(self.world):
            if d == self.rank:
                out.append(torch.zeros(0, dtype=torch.uint8))
                continue
            sel, cap = msgs[d], self.cap[d]
            if sel.size > cap:
                self.error |= ERR_WINDOW
            hdr = np.zeros(HDR_WORDS, np.uint32)
            hdr[0], hdr[1], hdr[2] = sel.size, self.error, n_nat
            hdr[4:12] = e_row
            pv = np.zeros((cap, 4), np.float32)
            og = np.zeros(cap, np.uint32)
            k = min(sel.size, cap)
            pv[:k, 0:2], pv[:k, 2:4], og[:k] = self.pos[sel[:k]], self.vel[sel[:k]], origin[sel[:k]]
            out.append(torch.from_numpy(np.concatenate([hdr.view(np.uint8), pv.view(np.uint8).ravel(), og.view(np.uint8)])))
        return out

    # ---- 3.-5. received records, counts, sort by (key, origin), tables ----
    def recv_sort(self, inbox):
        n_nat = self.pos.shape[0]
        pos, vel, org, n_all = [self.pos], [self.vel], [self._nat_origin], self.n_all.copy()
        for s_ in range(self.world):
            if s_ == self.rank:
                e = self.e_sent
            else:
                raw = inbox[s_].numpy()
                cap = self.cap[s_]
                hdr = raw[: HDR_WORDS * 4].view(np.uint32)
                if hdr[0] > cap:
                    self.error |= ERR_WINDOW
                self.error |= int(hdr[1])
                k = min(int(hdr[0]), cap)
                pv = raw[HDR_WORDS * 4: HDR_WORDS * 4 + cap * 16].view(np.float32).reshape(cap, 4)
                og = raw[HDR_WORDS * 4 + cap * 16:].view(np.uint32)
                pos.append(pv[:k, 0:2]); vel.append(pv[:k, 2:4]); org.append(og[:k])
                e = hdr[4:12]
            for r in range(self.world):
                n_all[r] += int(e[r]); n_all[s_] -= int(e[r])
        pos, vel, org = np.concatenate(pos), np.concatenate(vel), np.concatenate(org)
        if self.capacity is not None and pos.shape[0] > self.capacity:
            self.error |= ERR_CAPACITY
        self.n_all = n_all
        goff_new = int(n_all[: self.rank].sum())
        q = self._keys(pos, vel)
        row = (np.minimum(q["grid"], self.gw * self.gh - 1) // self.gw).astype(np.int64)
        keys = np.where(self.in_ext(self.rank, row), q["grid"], SENTINEL).astype(np.uint32)
        keys[:n_nat] = self._nat_keys
        keep = np.nonzero(keys != SENTINEL)[0]
        order = keep[np.lexsort((org[keep], keys[keep]))]  # in-cell order = by origin (the global index after the previous tick)
        ext = q[order]
        lo, hi = self.bounds[self.rank], self.bounds[self.rank + 1]
        rows = ext["grid"].astype(np.int64) // self.gw
        own_begin, own_end = int((rows < lo).sum()), int((rows < hi).sum())
        dens_begin, dens_end = int((rows < lo - 1).sum()), int((rows < hi + 1).sum())
        if own_end - own_begin != int(n_all[self.rank]):
            self.error |= ERR_COUNT
        # reference table rule with GLOBAL indices (compute.wgsl:45-56): every cell this rank sees, except global particle 0
        k = ext["grid"]
        first = np.nonzero(np.r_[True, k[1:] != k[:-1]])[0] if k.size else np.zeros(0, np.int64)
        gidx = goff_new + first - own_begin
        quirk = None
        if first.size and gidx[0] == 0:
            quirk = (int(k[0]), int(self.table[k[0]]))  # (lowest occupied cell, its stale GLOBAL entry)
            first, gidx = first[1:], gidx[1:]
        self.table[k[first]] = gidx.astype(np.uint32)
        self.ext, self.own, self.dens, self.goff_prev, self.goff, self.quirk = ext, (own_begin, own_end), (dens_begin, dens_end), self.goff, goff_new, quirk
        self.origin_explicit = None

    # ---- 6. density over owned + one ghost row per side, force over owned ----
    def density_force(self):
        ext, (ob, oe) = self.ext, self.own
        n = ext.shape[0]
        k = ext["grid"]
        local = np.full(self.gw * self.gh, 0xFFFFFFFF, np.uint32)  # empty cells read as empty
        first = np.nonzero(np.r_[True, k[1:] != k[:-1]])[0] if n else np.zeros(0, np.int64)
        local[k[first]] = first.astype(np.uint32)
        if self.quirk is not None:
            c0, stale = self.quirk
            local[c0] = min(max(stale - self.goff + ob, 0), 0xFFFFFFFF)
        u = O.make_uniform(self.s, self.ts, self.tick_count)
        u.particle_count = n
        if n:
            O.calculate_density(u, ext, local, 3)
            moved = O.move_particle_range(u, ext, local, self.texture, ob, oe, (self.goff - ob) & 0xFFFFFFFF)
        else:
            moved = ext
        self.sorted = moved[ob:oe].copy()
        self.pos, self.vel = self.sorted["position"].copy(), self.sorted["velocity"].copy()


def exchange(dist, rank, world, out, caps_in):
    """fixed-size windows to and from every other rank (what one ncclGroup of send/recv does on the device)."""
    inbox = [torch.zeros(0 if s == rank else HDR_WORDS * 4 + caps_in[s] * 20, dtype=torch.uint8) for s in range(world)]
    ops = []
    for r in range(world):
        if r == rank:
            continue
        ops.append(dist.P2POp(dist.isend, out[r], r))
        ops.append(dist.P2POp(dist.irecv, inbox[r], r))
    for w in dist.batch_isend_irecv(ops):
        w.wait()
    return inbox

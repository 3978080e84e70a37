"""CPU stand-in for gpu_fluid_simulation_b200.slab.SlabRank, built on the oracle (test infrastructure only).

It exposes the same phase methods, so the product's DistDriver — the code that decides who sends what to whom, in
which order, with which counts — can run under torch.distributed/gloo with world_size > 1 on a machine without a
GPU. The numerical passes are the oracle's functions applied to [halo | band | halo] arrays."""
import contextlib

import numpy as np
import torch

from gpu_fluid_simulation_b200.slab import NO_QUIRK, row_bounds
from oracle import oracle as O


class CpuSlabRank:
    def __init__(self, settings, rank, world):
        self.settings, self.rank, self.world = settings, rank, world
        self.gw, self.gh = O.grid_dims(settings)
        self.bounds = row_bounds(self.gh, world)
        self.p = np.zeros(0, dtype=O.PARTICLE_DTYPE)  # the band, global array order
        self.table = np.zeros(self.gw * self.gh, dtype=np.uint32)  # reference table, GLOBAL indices, owned cells
        tw, th = settings.texture_size
        self.texture = np.zeros((th, tw, 2), dtype=np.float32)
        self.tick_count = 0
        self.device = torch.device("cpu")
        self.halo_below = self.halo_above = 0

    def stream_ctx(self):
        return contextlib.nullcontext()

    def owns_rows(self):
        return self.bounds[self.rank], self.bounds[self.rank + 1]

    @property
    def n(self):
        return int(self.p.shape[0])

    def upload_band(self, particles):
        self.p = np.array(particles, dtype=O.PARTICLE_DTYPE, copy=True)

    def download_band(self):
        return self.p.copy()

    def sync(self):
        pass

    def _uniform(self, count):
        u = O.make_uniform(self.settings, self.ts, self.tick_count)
        u.particle_count = count
        return u

    def _keys(self, arr):
        tmp = arr.copy()
        u = self._uniform(tmp.shape[0])
        O.predict_next_position(u, tmp)
        O.create_spatial_lookup(u, tmp)
        return tmp

    # ---- phase 1 ----
    def begin_tick(self, ts):
        self.ts = ts
        self.tick_count += 1
        keyed = self._keys(self.p)
        rows = keyed["grid"] // self.gw
        lo, hi = self.owns_rows()
        self._emig = np.nonzero((rows < lo) | (rows >= hi))[0]
        dest = np.searchsorted(np.asarray(self.bounds[1:-1]), rows[self._emig], side="right")
        order = np.lexsort((self._emig, dest))  # grouped by destination, source order inside a group
        self._emig, self._dest = self._emig[order], dest[order]

    def send_counts(self):
        return torch.from_numpy(np.bincount(self._dest, minlength=self.world).astype(np.int64))

    def pack(self, send_counts_host):
        assert int(sum(send_counts_host)) == self._emig.size
        pay = np.concatenate([self.p["position"][self._emig], self.p["velocity"][self._emig]], axis=1).astype(np.float32)
        return torch.from_numpy(np.ascontiguousarray(pay).reshape(-1, 4))

    # ---- phase 2 ----
    def unpack(self, recv, recv_counts_host):
        r = recv.numpy().reshape(-1, 4)
        nb = int(sum(recv_counts_host[: self.rank]))
        na = int(sum(recv_counts_host[self.rank + 1:]))
        keep = np.ones(self.p.shape[0], bool)
        keep[self._emig] = False

        def mk(rows):
            q = np.zeros(rows.shape[0], dtype=O.PARTICLE_DTYPE)
            q["position"], q["velocity"] = rows[:, 0:2], rows[:, 2:4]
            return q
        self.p = np.concatenate([mk(r[:nb]), self.p[keep], mk(r[nb:nb + na])])

    # ---- phases 3+4 ----
    def sort(self):
        keyed = self._keys(self.p)
        O.sort_stable(keyed)
        self.p = keyed  # position / velocity untouched, predicted_position and grid of this tick, in key order
        lo, hi = self.owns_rows()
        rows = self.p["grid"] // self.gw
        self._first = int((rows == lo).sum()) if lo < hi else 0
        self._last = int((rows == hi - 1).sum()) if lo < hi else 0
        if lo == hi - 1:
            self._first = self._last = self.n

    def meta_tensor(self):
        k0 = int(self.p["grid"][0]) if self.n else NO_QUIRK
        stale = int(self.table[k0]) if self.n else 0
        return torch.tensor([self.n, self._first, self._last, k0, stale], dtype=torch.int64)

    # ---- phase 5 ----
    def halo_out(self, first_count, last_count):
        f, l = int(first_count), int(last_count)
        assert (f, l) == (self._first, self._last)
        self._out = [np.ascontiguousarray(self.p[k][:f]) for k in ("predicted_position", "velocity")] + \
                    [np.ascontiguousarray(self.p[k][self.n - l:]) for k in ("predicted_position", "velocity")]
        t = [torch.from_numpy(a) for a in self._out]
        return [t[0], t[1]], [t[2], t[3]]

    def halo_in(self, below_count, above_count):
        self.halo_below, self.halo_above = int(below_count), int(above_count)
        self._in = [np.zeros((self.halo_below, 2), np.float32), np.zeros((self.halo_below, 2), np.float32),
                    np.zeros((self.halo_above, 2), np.float32), np.zeros((self.halo_above, 2), np.float32)]
        t = [torch.from_numpy(a) for a in self._in]
        return [t[0], t[1]], [t[2], t[3]]

    def _halo_particles(self, pred, vel):
        q = np.zeros(pred.shape[0], dtype=O.PARTICLE_DTYPE)
        q["predicted_position"], q["velocity"] = pred, vel
        u = self._uniform(q.shape[0])
        O.create_spatial_lookup(u, q)
        return q

    # ---- phases 6+7 ----
    def cells_density(self, global_offset, quirk_c0, quirk_stale_global):
        self.global_offset = int(global_offset)
        hb, ha, n = self.halo_below, self.halo_above, self.n
        self.ext = np.concatenate([self._halo_particles(self._in[0], self._in[1]), self.p, self._halo_particles(self._in[2], self._in[3])])
        keys = self.ext["grid"]
        assert (np.diff(keys.astype(np.int64)) >= 0).all(), "[halo | band | halo] must be in key order"
        # local walk table (local indices): clean starts, plus the stale entry of the domain's lowest occupied cell
        self.local_table = np.zeros(self.gw * self.gh, dtype=np.uint32)
        first = np.nonzero(np.r_[True, keys[1:] != keys[:-1]])[0] if keys.size else np.zeros(0, np.int64)
        self.local_table[keys[first]] = first.astype(np.uint32)
        # cells that are empty here must read as empty: point them past the end
        occupied = np.zeros(self.gw * self.gh, bool)
        occupied[keys] = True
        self.local_table[~occupied] = 0xFFFFFFFF
        if quirk_c0 != NO_QUIRK and occupied[quirk_c0]:
            s = int(quirk_stale_global) - self.global_offset + hb
            self.local_table[quirk_c0] = min(max(s, 0), 0xFFFFFFFF)
        # reference-compatible table: owned particles write their GLOBAL index (global particle 0 writes nothing)
        for i in first:
            if hb <= i < hb + n:
                g = int(i) - hb + self.global_offset
                if g > 0:
                    self.table[keys[i]] = g
        u = self._uniform(self.ext.shape[0])
        O.calculate_density(u, self.ext, self.local_table, 3)
        self.p = self.ext[hb:hb + n].copy()

    def rho_out(self):
        self._rho_o = [np.ascontiguousarray(self.p["density"][: self._first]), np.ascontiguousarray(self.p["density"][self.n - self._last:])]
        return torch.from_numpy(self._rho_o[0]), torch.from_numpy(self._rho_o[1])

    def rho_in(self):
        self._rho_i = [np.zeros(self.halo_below, np.float32), np.zeros(self.halo_above, np.float32)]
        return torch.from_numpy(self._rho_i[0]), torch.from_numpy(self._rho_i[1])

    def force(self):
        hb, n = self.halo_below, self.n
        if getattr(self, "_rwi", None) is not None:  # device-resolved protocol: rho arrived in fixed windows
            self._rho_i = [self._rwi[0][self.pad - hb:].copy(), self._rwi[1][:self.halo_above].copy()]
            self._rwi = None
        self.ext["density"][:hb] = self._rho_i[0]
        self.ext["density"][hb + n:] = self._rho_i[1]
        u = self._uniform(self.ext.shape[0])
        out = O.move_particle_range(u, self.ext, self.local_table, self.texture, hb, hb + n, self.global_offset - hb)
        self.p = out[hb:hb + n].copy()
        self._first_saved = self._first

    # ---- device-resolved protocol (fixed-size windows, metadata consumed without a driver-side read) ----------------
    pad = 4096

    def meta_raw(self):
        m = self.meta_tensor().numpy()
        out = np.zeros(8, dtype=np.int64)
        out[:5] = m
        return torch.from_numpy(out.astype(np.uint32).view(np.int32).copy())

    @staticmethod
    def _window(arr, pad, right_aligned):
        """first / last `pad` rows of arr as a fixed-size window (zero filled where the band is shorter)."""
        w = np.zeros((pad,) + arr.shape[1:], dtype=arr.dtype)
        k = min(pad, arr.shape[0])
        if right_aligned:
            w[pad - k:] = arr[arr.shape[0] - k:]
        else:
            w[:k] = arr[:k]
        return w

    def halo_windows(self):
        pad = self.pad
        self._wout = [self._window(self.p[k], pad, False) for k in ("predicted_position", "velocity")] + \
                     [self._window(self.p[k], pad, True) for k in ("predicted_position", "velocity")]
        self._win = [np.zeros((pad, 2), np.float32) for _ in range(4)]
        t, u = [torch.from_numpy(a) for a in self._wout], [torch.from_numpy(a) for a in self._win]
        return [t[0], t[1]], [t[2], t[3]], [u[0], u[1]], [u[2], u[3]]

    def cells_density_dev(self, allmeta):
        m = allmeta.numpy().view(np.uint32).reshape(self.world, 8).astype(np.int64)
        me = self.rank
        hb = int(m[me - 1, 2]) if me > 0 else 0
        ha = int(m[me + 1, 1]) if me + 1 < self.world else 0
        assert hb <= self.pad and ha <= self.pad
        prefix = int(m[:me, 0].sum())
        c0, stale = NO_QUIRK, 0
        for r in range(self.world):
            if m[r, 0] > 0:
                c0, stale = int(m[r, 3]), int(m[r, 4])
                break
        self.halo_below, self.halo_above = hb, ha
        pad = self.pad
        self._in = [self._win[0][pad - hb:], self._win[1][pad - hb:], self._win[2][:ha], self._win[3][:ha]]
        self.cells_density(prefix, c0, stale)

    def rho_windows(self):
        pad = self.pad
        self._rwo = [self._window(self.p["density"], pad, False), self._window(self.p["density"], pad, True)]
        self._rwi = [np.zeros(pad, np.float32), np.zeros(pad, np.float32)]
        return (torch.from_numpy(self._rwo[0]), torch.from_numpy(self._rwo[1]), torch.from_numpy(self._rwi[0]),
                torch.from_numpy(self._rwi[1]))

    def error_row(self):
        return torch.zeros(self.world, dtype=torch.int64)

    def raise_on_error_bits(self, bits):
        assert bits == 0

    def close(self):
        pass

"""The multi-GPU fluid step: host mirror of the C ABI's fluid_multi_* block.

The whole step — kernels and the tick's one exchange between the bands — runs inside libfluid_b200 (csrc/multi.inl);
this module only marshals arguments. `MultiFluidSimulation.new` drives every rank from this process (what a Rust
`FluidSimulation::new_multi` would bind); `MultiFluidSimulation.new_rank` is one rank of a one-process-per-GPU job
(bench.py under torchrun), where the only thing the launcher contributes is handing rank 0's id bytes to the others.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import PARTICLE_DTYPE
from .simulation import SimulationSettings, TickSettings

TRANSPORT_AUTO, TRANSPORT_NCCL, TRANSPORT_LOCAL, TRANSPORT_PEER = 0, 1, 2, 3
TRANSPORT_NAMES = {0: "auto", 1: "nccl", 2: "local", 3: "peer (CUDA IPC pull over NVLink)"}
MATH_EXACT, MATH_FAST = 0, 1


def _options(row_bounds=None, capacity=0, window_near=0, window_far=0, transport=TRANSPORT_AUTO):
    o = _lib.MultiOptions()
    if row_bounds is not None:
        for i, v in enumerate(row_bounds):
            o.row_bounds[i] = int(v)
    o.capacity, o.window_near, o.window_far, o.transport = int(capacity), int(window_near), int(window_far), int(transport)
    return o


def unique_id() -> bytes:
    """the job's NCCL id (rank 0 makes it; every rank passes the same bytes to new_rank)."""
    buf = (C.c_uint8 * 128)()
    _lib.check(_lib.load().fluid_multi_unique_id(buf))
    return bytes(buf)


def balanced_bounds(settings: SimulationSettings, particles: np.ndarray, world: int):
    """row bounds giving every rank about the same number of particles (dam break)."""
    out = (C.c_uint32 * (world + 1))()
    p = np.ascontiguousarray(particles, dtype=PARTICLE_DTYPE)
    cs = settings.to_c()
    _lib.check(_lib.load().fluid_multi_balanced_bounds(C.byref(cs), _lib.ptr(p), p.shape[0], world, out))
    return list(out)


class MultiFluidSimulation:
    """reference: FluidSimulation (src/simulation.rs:10-37) over several GPUs; same `tick`, same `tick_count`."""

    def __init__(self, handle, settings, lib):
        self._h, self.settings, self._lib = handle, settings, lib

    @classmethod
    def new(cls, devices, settings: SimulationSettings, **opt):
        lib = _lib.load()
        h = C.c_void_p()
        cs = settings.to_c()
        devs = (C.c_int * len(devices))(*[int(d) for d in devices])
        o = _options(**opt)
        _lib.check(lib.fluid_multi_create(C.byref(cs), len(devices), devs, C.byref(o), C.byref(h)))
        return cls(h, settings, lib)

    @classmethod
    def new_rank(cls, rank, world, id_bytes, device, settings: SimulationSettings, **opt):
        lib = _lib.load()
        h = C.c_void_p()
        cs = settings.to_c()
        o = _options(**opt)
        idb = (C.c_uint8 * 128)(*id_bytes) if id_bytes is not None else None
        _lib.check(lib.fluid_multi_create_rank(C.byref(cs), rank, world, idb, device, C.byref(o), C.byref(h)))
        return cls(h, settings, lib)

    # ---- stepping ---------------------------------------------------------------------------------------------
    def tick(self, ts: TickSettings):
        c = ts.to_c()
        _lib.check(self._lib.fluid_multi_tick(self._h, C.byref(c)))

    def tick_n(self, ts: TickSettings, n: int):
        c = ts.to_c()
        _lib.check(self._lib.fluid_multi_tick_n(self._h, C.byref(c), int(n)))

    def sync(self):
        _lib.check(self._lib.fluid_multi_sync(self._h))

    @property
    def tick_count(self):
        return self._lib.fluid_multi_tick_count(self._h)

    @tick_count.setter
    def tick_count(self, v):
        _lib.check(self._lib.fluid_multi_set_tick_count(self._h, int(v)))

    def set_math_mode(self, mode):
        _lib.check(self._lib.fluid_multi_set_math_mode(self._h, int(mode)))

    @property
    def world(self):
        return self._lib.fluid_multi_world(self._h)

    @property
    def local_bands(self):
        return self._lib.fluid_multi_local_bands(self._h)

    @property
    def transport(self):
        return self._lib.fluid_multi_transport(self._h)

    @property
    def row_bounds(self):
        out = (C.c_uint32 * (self.world + 1))()
        _lib.check(self._lib.fluid_multi_row_bounds(self._h, out))
        return list(out)

    @property
    def grid(self):
        s = self._lib.fluid_multi_band_sim(self._h, 0)
        return self._lib.fluid_sim_grid_w(s), self._lib.fluid_sim_grid_h(s)

    # ---- state ------------------------------------------------------------------------------------------------
    def upload_particles(self, particles: np.ndarray):
        p = np.ascontiguousarray(particles, dtype=PARTICLE_DTYPE)
        _lib.check(self._lib.fluid_multi_upload_particles(self._h, _lib.ptr(p), p.shape[0]))

    def download_particles(self) -> np.ndarray:
        out = np.empty(self.settings.particle_count, dtype=PARTICLE_DTYPE)
        _lib.check(self._lib.fluid_multi_download_particles(self._h, _lib.ptr(out)))
        return out

    def upload_band(self, particles: np.ndarray, origins=None):
        p = np.ascontiguousarray(particles, dtype=PARTICLE_DTYPE)
        o = None if origins is None else np.ascontiguousarray(origins, dtype=np.uint32)
        _lib.check(self._lib.fluid_multi_upload_band(self._h, _lib.ptr(p) if p.shape[0] else None, p.shape[0], _lib.ptr(o)))

    def download_band(self, local_band=0, out=None):
        """(particles of the band in key order, its offset in the global key order)."""
        if out is None:
            counts = self.particle_counts()
            bounds_rank = self._rank_of_local(local_band)
            out = np.empty(int(counts[bounds_rank]), dtype=PARTICLE_DTYPE)
        n, goff = C.c_uint32(), C.c_uint32()
        _lib.check(self._lib.fluid_multi_download_band(self._h, local_band, _lib.ptr(out) if out.shape[0] else None, out.shape[0],
                                                       C.byref(n), C.byref(goff)))
        return out[: n.value], goff.value

    def upload_band_async(self, particles):
        """pinned AoS source; collective; complete after sync_transfers()."""
        n = int(particles.shape[0])
        _lib.check(self._lib.fluid_multi_upload_band_async(self._h, _lib.ptr(particles) if n else None, n))

    def download_band_async(self, out, info):
        """out: pinned AoS array (its length = the copy's capacity); info: pinned uint32[2] <- (size, global offset)."""
        _lib.check(self._lib.fluid_multi_download_band_async(self._h, _lib.ptr(out), int(out.shape[0]), _lib.ptr(info)))

    def sync_transfers(self):
        _lib.check(self._lib.fluid_multi_sync_transfers(self._h))

    def _rank_of_local(self, local_band):
        return local_band if self.local_bands == self.world else getattr(self, "rank", 0)

    def particle_counts(self):
        out = (C.c_uint32 * self.world)()
        _lib.check(self._lib.fluid_multi_particle_counts(self._h, out))
        return list(out)

    def download_start_indices(self) -> np.ndarray:
        gw, gh = self.grid
        out = np.empty(gw * gh, dtype=np.uint32)
        _lib.check(self._lib.fluid_multi_download_start_indices(self._h, _lib.ptr(out)))
        return out

    def upload_start_indices(self, table: np.ndarray):
        t = np.ascontiguousarray(table, dtype=np.uint32)
        _lib.check(self._lib.fluid_multi_upload_start_indices(self._h, _lib.ptr(t)))

    def upload_force_field(self, field: np.ndarray):
        f = np.ascontiguousarray(field, dtype=np.float32)
        _lib.check(self._lib.fluid_multi_upload_force_field(self._h, _lib.ptr(f)))

    # ---- measurement --------------------------------------------------------------------------------------------
    def enable_timing(self, on=True):
        _lib.check(self._lib.fluid_multi_enable_timing(self._h, 1 if on else 0))

    def last_timing(self, local_band=0):
        names = (C.c_char_p * 16)()
        ms = (C.c_float * 16)()
        k = self._lib.fluid_multi_last_timing(self._h, local_band, names, ms, 16)
        if k < 0:
            _lib.check(k)
        return [(names[i].decode(), float(ms[i])) for i in range(k)]

    @property
    def launch_count(self):
        return int(self._lib.fluid_multi_launch_count(self._h))

    def stream(self, local_band=0):
        return self._lib.fluid_multi_stream(self._h, local_band)

    def close(self):
        if self._h:
            self._lib.fluid_multi_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

"""Host-side mirror of the reference's simulation API (reference src/simulation.rs).

Same names, argument meaning and error behaviour as the Rust types, on top of the C ABI:

  SimulationSettings   src/simulation.rs:95-104
  TickSettings         src/simulation.rs:107-122   (defaults: src/renderer.rs:374-388)
  FluidSimulation      src/simulation.rs:10-37     new() :139, tick() :459, pub tick :12,
                                                   force_field_texture() :562

The wgpu bind-group getters (:542-560) have no meaning without a graphics API; their headless
stand-ins are download_particles / start_indices / device_view.
"""
import ctypes as C
from dataclasses import dataclass, field
from typing import Tuple

import numpy as np

from . import _lib
from ._lib import PARTICLE_DTYPE, FluidError


@dataclass
class SimulationSettings:
    """reference: SimulationSettings, src/simulation.rs:95-104; defaults = the default scene, src/main.rs:48-54."""
    particle_count: int = 100_000
    particle_spacing: float = 0.1
    smoothing_radius: float = 0.2
    size: Tuple[float, float] = (53.0, 53.0)
    texture_size: Tuple[int, int] = (1024, 1024)

    def to_c(self):
        s = _lib.SimSettings()
        s.particle_count = int(self.particle_count)
        s.particle_spacing = float(self.particle_spacing)
        s.smoothing_radius = float(self.smoothing_radius)
        s.size[0], s.size[1] = float(self.size[0]), float(self.size[1])
        s.texture_size[0], s.texture_size[1] = int(self.texture_size[0]), int(self.texture_size[1])
        return s


@dataclass
class TickSettings:
    """reference: TickSettings, src/simulation.rs:107-122; default values from src/renderer.rs:374-388."""
    delta: float = 1.0 / 120.0
    gravity: Tuple[float, float] = (0.0, 0.0)
    mass: float = 1.0
    pressure_constant: float = 50.0
    rest_density: float = 0.0
    damping_factor: float = 0.1
    viscosity_coefficient: float = 25.0
    surface_tension_treshold: float = 0.1
    surface_tension_coefficient: float = 35.0
    mouse_force_radius: float = 5.0
    mouse_force_power: float = 150.0
    mouse_pos: Tuple[float, float] = (0.0, 0.0)
    mouse_state: int = 0

    def to_c(self):
        t = _lib.TickSettingsC()
        t.delta = self.delta
        t.gravity[0], t.gravity[1] = self.gravity
        t.mass = self.mass
        t.pressure_constant = self.pressure_constant
        t.rest_density = self.rest_density
        t.damping_factor = self.damping_factor
        t.viscosity_coefficient = self.viscosity_coefficient
        t.surface_tension_treshold = self.surface_tension_treshold
        t.surface_tension_coefficient = self.surface_tension_coefficient
        t.mouse_force_radius = self.mouse_force_radius
        t.mouse_force_power = self.mouse_force_power
        t.mouse_pos[0], t.mouse_pos[1] = self.mouse_pos
        t.mouse_state = int(self.mouse_state)
        return t


class FluidSimulation:
    """reference: FluidSimulation, src/simulation.rs:10-37.

    `FluidSimulation.new(device, settings)` mirrors `FluidSimulation::new(&Device, SimulationSettings)`
    (:139); here `device` is a CUDA ordinal. `tick(settings)` mirrors `tick(&Queue, &mut Encoder,
    TickSettings)` (:459): it only enqueues work, `sync()` is the reference's `device.poll(Wait)`.
    particle_count <= 1 raises (the reference panics, :323-324).
    """

    def __init__(self, device: int, settings: SimulationSettings):
        self._lib = _lib.load()
        self.settings = settings
        self._h = C.c_void_p()
        cs = settings.to_c()
        _lib.check(self._lib.fluid_sim_create(C.byref(cs), int(device), C.byref(self._h)))
        self.device = int(device)

    @classmethod
    def new(cls, device: int, settings: SimulationSettings) -> "FluidSimulation":
        return cls(device, settings)

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.fluid_sim_destroy(self._h)
            self._h = C.c_void_p()

    def restart(self):
        """reference: Renderer::restart_simulation (src/renderer.rs:870-872) — a fresh FluidSimulation::new with the
        same SimulationSettings: spawn grid, zero velocities, tick 0, start table and field re-initialised."""
        self.close()
        cs = self.settings.to_c()
        _lib.check(self._lib.fluid_sim_create(C.byref(cs), self.device, C.byref(self._h)))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- stepping ---------------------------------------------------------------------------
    @property
    def tick_count(self) -> int:
        """reference: `pub tick: u32` (src/simulation.rs:12)."""
        return int(self._lib.fluid_sim_tick_count(self._h))

    @tick_count.setter
    def tick_count(self, v: int):
        _lib.check(self._lib.fluid_sim_set_tick_count(self._h, int(v)))

    def tick(self, settings: TickSettings):
        ts = settings.to_c()
        _lib.check(self._lib.fluid_sim_tick(self._h, C.byref(ts)))

    def tick_n(self, settings: TickSettings, n: int):
        ts = settings.to_c()
        _lib.check(self._lib.fluid_sim_tick_n(self._h, C.byref(ts), int(n)))

    def tick_hash_sort_only(self, settings: TickSettings):
        ts = settings.to_c()
        _lib.check(self._lib.fluid_sim_tick_hash_sort_only(self._h, C.byref(ts)))

    MATH_EXACT, MATH_FAST = 0, 1

    @property
    def math_mode(self) -> int:
        return int(self._lib.fluid_sim_get_math_mode(self._h))

    @math_mode.setter
    def math_mode(self, mode: int):
        _lib.check(self._lib.fluid_sim_set_math_mode(self._h, int(mode)))

    SORT_AUTO, SORT_RADIX, SORT_COUNTING = 0, 1, 2

    @property
    def sort_mode(self) -> int:
        return int(self._lib.fluid_sim_get_sort_mode(self._h))

    @sort_mode.setter
    def sort_mode(self, mode: int):
        _lib.check(self._lib.fluid_sim_set_sort_mode(self._h, int(mode)))

    @property
    def last_sort_path(self) -> int:
        """SORT_RADIX or SORT_COUNTING: what the most recently enqueued tick used."""
        return int(self._lib.fluid_sim_last_sort_path(self._h))

    def sync(self):
        _lib.check(self._lib.fluid_sim_sync(self._h))

    # -- geometry ---------------------------------------------------------------------------
    @property
    def particle_count(self) -> int:
        return int(self._lib.fluid_sim_particle_count(self._h))

    @property
    def grid_w(self) -> int:
        return int(self._lib.fluid_sim_grid_w(self._h))

    @property
    def grid_h(self) -> int:
        return int(self._lib.fluid_sim_grid_h(self._h))

    # -- state exposure (headless stand-ins for simulation_bg / force_field_texture) ---------
    def download_particles(self, out=None) -> np.ndarray:
        n = self.particle_count
        if out is None:
            out = np.empty(n, dtype=PARTICLE_DTYPE)
        _lib.check(self._lib.fluid_sim_download_particles(self._h, _lib.ptr(out)))
        return out

    def upload_particles(self, particles):
        if isinstance(particles, np.ndarray):
            assert particles.dtype == PARTICLE_DTYPE and particles.shape == (self.particle_count,)
            particles = np.ascontiguousarray(particles)
        _lib.check(self._lib.fluid_sim_upload_particles(self._h, _lib.ptr(particles)))

    # pipelined transfers: copies on the handle's own streams, upload of the next state overlapping the download of
    # the previous result; buffers should be pinned and stay untouched until sync_transfers()
    def upload_particles_async(self, particles):
        if isinstance(particles, np.ndarray):
            assert particles.dtype == PARTICLE_DTYPE and particles.shape == (self.particle_count,)
            assert particles.flags["C_CONTIGUOUS"]
        _lib.check(self._lib.fluid_sim_upload_particles_async(self._h, _lib.ptr(particles)))

    def download_particles_async(self, out):
        _lib.check(self._lib.fluid_sim_download_particles_async(self._h, _lib.ptr(out)))

    def state_changed(self):
        """call after writing position / velocity through the device pointers of device_view()"""
        _lib.check(self._lib.fluid_sim_state_changed(self._h))

    def sync_transfers(self):
        _lib.check(self._lib.fluid_sim_sync_transfers(self._h))

    def download_start_indices(self) -> np.ndarray:
        out = np.empty(self.grid_w * self.grid_h, dtype=np.uint32)
        _lib.check(self._lib.fluid_sim_download_start_indices(self._h, _lib.ptr(out)))
        return out

    def upload_start_indices(self, table: np.ndarray):
        table = np.ascontiguousarray(table, dtype=np.uint32)
        assert table.size == self.grid_w * self.grid_h
        _lib.check(self._lib.fluid_sim_upload_start_indices(self._h, _lib.ptr(table)))

    def upload_force_field(self, field_xy: np.ndarray):
        """reference: queue.write_buffer(simulation.force_field_texture(), ..), src/renderer.rs:498-502."""
        f = np.ascontiguousarray(field_xy, dtype=np.float32)
        assert f.size == 2 * self.settings.texture_size[0] * self.settings.texture_size[1]
        _lib.check(self._lib.fluid_sim_upload_force_field(self._h, _lib.ptr(f)))

    def download_force_field(self) -> np.ndarray:
        tw, th = self.settings.texture_size
        out = np.empty((th, tw, 2), dtype=np.float32)
        _lib.check(self._lib.fluid_sim_download_force_field(self._h, _lib.ptr(out)))
        return out

    def uniform(self) -> "_lib.Uniform":
        u = _lib.Uniform()
        _lib.check(self._lib.fluid_sim_get_uniform(self._h, C.byref(u)))
        return u

    # -- parity taps -------------------------------------------------------------------------
    def download_unsorted_keys(self) -> np.ndarray:
        out = np.empty(self.particle_count, dtype=np.uint32)
        _lib.check(self._lib.fluid_sim_download_unsorted_keys(self._h, _lib.ptr(out)))
        return out

    def download_permutation(self) -> np.ndarray:
        out = np.empty(self.particle_count, dtype=np.uint32)
        _lib.check(self._lib.fluid_sim_download_permutation(self._h, _lib.ptr(out)))
        return out

    def enable_debug_force(self, enable=True):
        _lib.check(self._lib.fluid_sim_enable_debug_force(self._h, int(bool(enable))))

    def download_debug_force(self):
        n = self.particle_count
        accel = np.empty((n, 2), dtype=np.float32)
        absum = np.empty((n, 2), dtype=np.float32)
        _lib.check(self._lib.fluid_sim_download_debug_force(self._h, _lib.ptr(accel), _lib.ptr(absum)))
        return accel, absum

    # -- device-side access / instrumentation -------------------------------------------------
    def device_view(self) -> "_lib.DeviceView":
        v = _lib.DeviceView()
        _lib.check(self._lib.fluid_sim_device_view(self._h, C.byref(v)))
        return v

    @property
    def stream(self) -> int:
        return int(self._lib.fluid_sim_stream(self._h) or 0)

    def enable_timing(self, enable=True):
        _lib.check(self._lib.fluid_sim_enable_timing(self._h, int(bool(enable))))

    def last_timing(self):
        names = (C.c_char_p * 16)()
        ms = (C.c_float * 16)()
        k = self._lib.fluid_sim_last_timing(self._h, names, ms, 16)
        if k < 0:
            _lib.check(k)
        return [(names[i].decode(), float(ms[i])) for i in range(k)]

    @property
    def launch_count(self) -> int:
        return int(self._lib.fluid_sim_launch_count(self._h))

    def reserve(self, capacity: int):
        _lib.check(self._lib.fluid_sim_reserve(self._h, int(capacity)))

    def set_particle_count(self, n: int):
        _lib.check(self._lib.fluid_sim_set_particle_count(self._h, int(n)))

    # -- jump flood ---------------------------------------------------------------------------
    def jfa_from_particles(self, w: int, h: int, want_seeds=True, want_field=True, into_sim=False):
        seeds = np.empty((h, w, 2), dtype=np.uint32) if want_seeds else None
        fld = np.empty((h, w, 2), dtype=np.float32) if want_field else None
        ms = C.c_float(0)
        _lib.check(self._lib.fluid_sim_jfa_from_particles(self._h, int(w), int(h), _lib.ptr(seeds), _lib.ptr(fld),
                                                        int(bool(into_sim)), C.byref(ms)))
        return seeds, fld, float(ms.value)

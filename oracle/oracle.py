"""ctypes binding of oracle/liboracle.so — the CPU restatement of the reference's fluid step.

TEST INFRASTRUCTURE: the checker for tests/, __graft_entry__.smoke() and bench.py's cpu_baseline and
`--impl reference` legs. Nothing in gpu_fluid_simulation_b200/ imports this module.
PARITY: the reference has no tests or golden vectors and cannot run here; this restatement is pinned instead by a second,
mechanical derivation — the reference's WGSL transpiled to C++ (oracle/transpile_wgsl.py, oracle/wgsl_ref.py) — against which
tests/test_wgsl_ref.py compares it bit for bit. The Rust host parts (uniform, lattice, CPU field) remain restated by hand.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")

PARTICLE_DTYPE = np.dtype([
    ("position", np.float32, 2), ("predicted_position", np.float32, 2), ("velocity", np.float32, 2),
    ("density", np.float32), ("grid", np.uint32)], align=False)


class Settings(C.Structure):
    _fields_ = [("particle_count", C.c_uint32), ("particle_spacing", C.c_float), ("smoothing_radius", C.c_float),
                ("size", C.c_float * 2), ("texture_size", C.c_uint32 * 2)]


class Tick(C.Structure):
    _fields_ = [("delta", C.c_float), ("gravity", C.c_float * 2), ("mass", C.c_float), ("pressure_constant", C.c_float),
                ("rest_density", C.c_float), ("damping_factor", C.c_float), ("viscosity_coefficient", C.c_float),
                ("surface_tension_treshold", C.c_float), ("surface_tension_coefficient", C.c_float),
                ("mouse_force_radius", C.c_float), ("mouse_force_power", C.c_float), ("mouse_pos", C.c_float * 2),
                ("mouse_state", C.c_int32)]


class Uniform(C.Structure):
    _fields_ = [("delta", C.c_float), ("particle_count", C.c_uint32), ("sqr_radius", C.c_float), ("frame_time", C.c_uint32),
                ("gravity", C.c_float * 2), ("bounds", C.c_float * 2), ("mouse_pos", C.c_float * 2),
                ("smoothing_radius", C.c_float), ("particle_mass", C.c_float), ("pressure_constant", C.c_float),
                ("rest_density", C.c_float), ("damping_factor", C.c_float), ("viscosity_coefficient", C.c_float),
                ("surface_tension_treshold", C.c_float), ("surface_tension_coefficient", C.c_float),
                ("poly6_kernel_volume", C.c_float), ("poly6_kernel_derivative", C.c_float),
                ("poly6_kernel_laplacian", C.c_float), ("spiky_kernel_derivative", C.c_float),
                ("viscosity_kernel", C.c_float), ("mouse_state", C.c_int32), ("mouse_force_radius", C.c_float),
                ("mouse_force_power", C.c_float), ("grid_w", C.c_uint32), ("grid_h", C.c_uint32),
                ("texture_size", C.c_float * 2)]


def build(force=False):
    srcs = [os.path.join(HERE, f) for f in ("fluid_oracle.cpp", "render_oracle.cpp")]
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(f) for f in srcs):
        subprocess.run(["make", "-C", HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return LIB_PATH


_lib = None
_vp = C.c_void_p


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.oracle_poly6_norm.restype = C.c_float
        L.oracle_poly6_norm.argtypes = [C.c_float]
        L.oracle_jfa_first_jump.restype = C.c_uint32
        L.oracle_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return None if a is None else _vp(a.ctypes.data)


def settings_c(s):
    """accepts the product's SimulationSettings dataclass (duck-typed) and returns the oracle's C struct."""
    c = Settings()
    c.particle_count = int(s.particle_count)
    c.particle_spacing = float(s.particle_spacing)
    c.smoothing_radius = float(s.smoothing_radius)
    c.size[0], c.size[1] = float(s.size[0]), float(s.size[1])
    c.texture_size[0], c.texture_size[1] = int(s.texture_size[0]), int(s.texture_size[1])
    return c


def tick_c(t):
    c = Tick()
    c.delta = t.delta
    c.gravity[0], c.gravity[1] = t.gravity
    c.mass = t.mass
    c.pressure_constant = t.pressure_constant
    c.rest_density = t.rest_density
    c.damping_factor = t.damping_factor
    c.viscosity_coefficient = t.viscosity_coefficient
    c.surface_tension_treshold = t.surface_tension_treshold
    c.surface_tension_coefficient = t.surface_tension_coefficient
    c.mouse_force_radius = t.mouse_force_radius
    c.mouse_force_power = t.mouse_force_power
    c.mouse_pos[0], c.mouse_pos[1] = t.mouse_pos
    c.mouse_state = int(t.mouse_state)
    return c


def grid_dims(s):
    gw, gh = C.c_uint32(), C.c_uint32()
    cs = settings_c(s)
    lib().oracle_grid_dims(C.byref(cs), C.byref(gw), C.byref(gh))
    return gw.value, gh.value


def init_lattice(s):
    out = np.zeros(int(s.particle_count), dtype=PARTICLE_DTYPE)
    cs = settings_c(s)
    lib().oracle_init_lattice(C.byref(cs), _p(out))
    return out


def make_uniform(s, t, tick):
    u = Uniform()
    cs, ct = settings_c(s), tick_c(t)
    lib().oracle_make_uniform(C.byref(cs), C.byref(ct), C.c_uint32(tick), C.byref(u))
    return u


def predict_next_position(u, p):
    lib().oracle_predict_next_position(C.byref(u), _p(p))


def create_spatial_lookup(u, p):
    lib().oracle_create_spatial_lookup(C.byref(u), _p(p))


def sort_bitonic(p):
    lib().oracle_sort_bitonic(_p(p), C.c_uint32(p.shape[0]))


def sort_stable(p):
    perm = np.empty(p.shape[0], dtype=np.uint32)
    lib().oracle_sort_stable(_p(p), C.c_uint32(p.shape[0]), _p(perm))
    return perm


def sort_counting(p, num_cells):
    lib().oracle_sort_counting(_p(p), C.c_uint32(p.shape[0]), C.c_uint32(num_cells))


def compute_start_indices(u, p, table):
    lib().oracle_compute_start_indices(C.byref(u), _p(p), _p(table))


def calculate_density(u, p, table, walk_radius=3):
    lib().oracle_calculate_density(C.byref(u), _p(p), _p(table), C.c_int(walk_radius))


def move_particle(u, p_in, table, texture, want_debug=False):
    out = np.empty_like(p_in)
    accel = np.zeros((p_in.shape[0], 2), np.float32) if want_debug else None
    absum = np.zeros((p_in.shape[0], 2), np.float32) if want_debug else None
    lib().oracle_move_particle(C.byref(u), _p(p_in), _p(out), _p(table), _p(texture), _p(accel), _p(absum))
    return (out, accel, absum) if want_debug else out


def move_particle_range(u, p_in, table, texture, begin, end, index_offset):
    """slab tests: move only [begin, end); RNG ids = array index + index_offset."""
    out = p_in.copy()
    lib().oracle_move_particle_range(C.byref(u), _p(p_in), _p(out), _p(table), _p(texture), C.c_uint32(begin), C.c_uint32(end),
                                     C.c_uint32(index_offset & 0xFFFFFFFF))
    return out


SORT_STABLE, SORT_BITONIC, SORT_COUNTING = 0, 1, 2


def tick(u, p, table, texture, sort_mode=SORT_STABLE, walk_radius=3, stop_after=0):
    """One full reference tick in place on (p, table). Returns per-phase milliseconds (6 floats)."""
    ms = np.zeros(6, dtype=np.float32)
    lib().oracle_tick(C.byref(u), _p(p), _p(table), _p(texture), C.c_int(sort_mode), C.c_int(walk_radius),
                      C.c_int(stop_after), _p(ms))
    return ms


class OracleSim:
    """Stateful convenience wrapper with the same shape as the reference's FluidSimulation
    (new / tick / pub tick), driving the oracle functions in the reference's dispatch order."""

    def __init__(self, settings, particles=None, sort_mode=SORT_STABLE, walk_radius=3):
        self.settings = settings
        self.gw, self.gh = grid_dims(settings)
        self.p = init_lattice(settings) if particles is None else np.array(particles, dtype=PARTICLE_DTYPE, copy=True)
        self.table = np.zeros(self.gw * self.gh, dtype=np.uint32)
        tw, th = settings.texture_size
        self.texture = np.zeros((th, tw, 2), dtype=np.float32)
        self.tick_count = 0
        self.sort_mode = sort_mode
        self.walk_radius = walk_radius
        self.last_uniform = None

    def tick(self, t, stop_after=0):
        self.tick_count += 1
        u = make_uniform(self.settings, t, self.tick_count)
        self.last_uniform = u
        return tick(u, self.p, self.table, self.texture, self.sort_mode, self.walk_radius, stop_after)


# ---- jump flood -------------------------------------------------------------------------------------
def jfa_pass(seeds, jump):
    seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
    h, w, _ = seeds.shape
    out = np.empty_like(seeds)
    lib().oracle_jfa_pass(_p(seeds), C.c_uint32(w), C.c_uint32(h), C.c_uint32(jump), _p(out))
    return out


def jfa_run(seeds):
    seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
    h, w, _ = seeds.shape
    out = np.empty_like(seeds)
    lib().oracle_jfa_run(_p(seeds), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def jfa_seed_from_mask(mask):
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    h, w = mask.shape
    out = np.empty((h, w, 2), dtype=np.uint32)
    lib().oracle_jfa_seed_from_mask(_p(mask), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def jfa_seed_from_particles(p, bounds, w, h):
    out = np.empty((h, w, 2), dtype=np.uint32)
    b = np.array(bounds, dtype=np.float32)
    lib().oracle_jfa_seed_from_particles(_p(p), C.c_uint32(p.shape[0]), _p(b), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def jfa_to_force_field(seeds):
    seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
    h, w, _ = seeds.shape
    out = np.empty((h, w, 2), dtype=np.float32)
    lib().oracle_jfa_to_force_field(_p(seeds), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def generate_smooth_gradient_field(mask, want_nearest=False):
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    h, w = mask.shape
    out = np.empty((h, w, 2), dtype=np.float32)
    near = np.empty((h, w, 2), dtype=np.uint32) if want_nearest else None
    lib().oracle_generate_smooth_gradient_field(_p(mask), C.c_uint32(w), C.c_uint32(h), _p(out), _p(near))
    return (out, near) if want_nearest else out


# ---- per-pixel shaders (render_oracle.cpp) ------------------------------------------------------------
OBJECT_DTYPE = np.dtype([("position", np.float32, 2), ("kind", np.uint32), ("pad", np.uint32), ("pad2", np.uint32, 4)], align=False)


def rasterize_objects(objects, inv_proj, w, h):
    """image_shader.wgsl fs_main -> u8 mask (0 inside an object, 255 elsewhere)."""
    objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    out = np.empty((h, w), dtype=np.uint8)
    lib().oracle_rasterize_objects(_p(objects), C.c_uint32(objects.shape[0]), _p(m), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def render_fluid(u, p, table, inv_proj, w, h):
    """fluid_shader.wgsl fs_main -> float32[h, w, 4] (the fragment output, before blending)."""
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    out = np.empty((h, w, 4), dtype=np.float32)
    lib().oracle_render_fluid(C.byref(u), _p(p), _p(table), _p(m), C.c_uint32(w), C.c_uint32(h), _p(out))
    return out


def surface_tension(u, p, table):
    """the reference's dormant calculate_surface_tension per particle -> float32[n, 2] (not applied to the state)."""
    out = np.empty((p.shape[0], 2), dtype=np.float32)
    lib().oracle_surface_tension(C.byref(u), _p(p), _p(table), _p(out))
    return out


def bruteforce_density_f64(u, p):
    out = np.empty(p.shape[0], dtype=np.float64)
    lib().oracle_bruteforce_density_f64(C.byref(u), _p(p), _p(out))
    return out


def bruteforce_accel_f64(u, p):
    out = np.empty((p.shape[0], 2), dtype=np.float64)
    lib().oracle_bruteforce_accel_f64(C.byref(u), _p(p), _p(out))
    return out

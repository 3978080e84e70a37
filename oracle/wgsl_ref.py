"""ctypes binding of oracle/_ref/libwgsl_ref.so — the reference's WGSL shaders, transpiled mechanically to C++ by
oracle/transpile_wgsl.py and driven by oracle/wgsl_ref_driver.cpp (TEST INFRASTRUCTURE: tests/ and bench.py's CPU legs
only). The library is generated from /root/reference at build time and is git-ignored; it travels to the GPU box as a
built artefact. available() is False where neither the library nor the reference sources exist."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_ref", "libwgsl_ref.so")
REFERENCE = "/root/reference"
_lib = None
_vp = C.c_void_p


def build(force=False):
    """needs the reference sources where they lie (this container); elsewhere the prebuilt library is used as is."""
    if not os.path.isdir(REFERENCE):
        return os.path.exists(LIB_PATH)
    srcs = [os.path.join(HERE, f) for f in ("transpile_wgsl.py", "wgsl_ref_driver.cpp", "wgsl_prelude.h")]
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < max(os.path.getmtime(f) for f in srcs):
        subprocess.run([sys.executable, os.path.join(HERE, "transpile_wgsl.py"), "--reference", REFERENCE], check=True, capture_output=True)
    return True


def available():
    return os.path.exists(LIB_PATH) or build()


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libwgsl_ref.so is missing and /root/reference is not here to generate it from")
        L = C.CDLL(LIB_PATH)
        L.wgslref_num_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a):
    return None if a is None else _vp(a.ctypes.data)


def tick(u, p, table, texture, stop_after=0):
    """FluidSimulation::tick for the uniform `u` (oracle.Uniform), in place on (p, table). Returns 6 phase times in ms."""
    ms = np.zeros(6, dtype=np.float32)
    tex = np.ascontiguousarray(texture, dtype=np.float32)
    lib().wgslref_tick(C.byref(u), _p(p), C.c_uint32(p.shape[0]), _p(table), C.c_uint32(table.shape[0]), _p(tex),
                       C.c_uint32(tex.size // 2), C.c_int(stop_after), _p(ms))
    return ms


def run_pass(name, u, p, table, texture):
    tex = np.ascontiguousarray(texture, dtype=np.float32)
    getattr(lib(), "wgslref_" + name)(C.byref(u), _p(p), C.c_uint32(p.shape[0]), _p(table), C.c_uint32(table.shape[0]), _p(tex),
                                      C.c_uint32(tex.size // 2))


def move_particle(u, p, table, texture):
    tex = np.ascontiguousarray(texture, dtype=np.float32)
    out = np.empty_like(p)
    lib().wgslref_move_particle(C.byref(u), _p(p), _p(out), C.c_uint32(p.shape[0]), _p(table), C.c_uint32(table.shape[0]), _p(tex),
                                C.c_uint32(tex.size // 2))
    return out


def sort(p):
    lib().wgslref_sort(_p(p), C.c_uint32(p.shape[0]))


def jfa_pass(seeds, jump):
    seeds = np.ascontiguousarray(seeds, dtype=np.uint32)
    h, w, _ = seeds.shape
    out = np.zeros_like(seeds)
    lib().wgslref_jfa_pass(_p(seeds), C.c_uint32(w), C.c_uint32(h), C.c_uint32(jump), _p(out))
    return out


class WgslRefSim:
    """same shape as oracle.OracleSim, stepping the transpiled shaders (bitonic sort, 7x7 density walk: the reference's own)."""

    def __init__(self, oracle_module, settings, particles=None):
        self.O = oracle_module
        self.settings = settings
        self.gw, self.gh = oracle_module.grid_dims(settings)
        self.p = oracle_module.init_lattice(settings) if particles is None else np.array(particles, dtype=oracle_module.PARTICLE_DTYPE, copy=True)
        self.table = np.zeros(self.gw * self.gh, dtype=np.uint32)
        tw, th = settings.texture_size
        self.texture = np.zeros((th, tw, 2), dtype=np.float32)
        self.tick_count = 0

    def tick(self, t, stop_after=0):
        self.tick_count += 1
        u = self.O.make_uniform(self.settings, t, self.tick_count)  # src/simulation.rs:470-497 (host code: restated, not transpiled)
        return tick(u, self.p, self.table, self.texture, stop_after)

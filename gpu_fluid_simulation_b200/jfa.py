"""Jump-flood distance field through the C ABI (reference: shaders/jump_flood.wgsl:17-65; host protocol
pinned in SURVEY.md §8 row a7; CPU routine it replaces: generate_smooth_gradient_field, src/main.rs:403-515)."""
import ctypes as C

import numpy as np

from . import _lib

INVALID = 32767


def _img(a, dtype, ch):
    a = np.ascontiguousarray(a, dtype=dtype)
    assert a.ndim == 3 and a.shape[2] == ch, a.shape
    return a


def jfa_run(seeds: np.ndarray, device: int = 0) -> np.ndarray:
    seeds = _img(seeds, np.uint32, 2)
    h, w, _ = seeds.shape
    out = np.empty_like(seeds)
    _lib.check(_lib.load().fluid_jfa_run(_lib.ptr(seeds), w, h, _lib.ptr(out), device))
    return out


JFA_STATIC_MAX_PIXELS = 1 << 20  # FLUID_JFA_STATIC_MAX_PIXELS (include/fluid_b200.h)


def jfa_run_static(seeds: np.ndarray, device: int = 0) -> np.ndarray:
    """fluid_jfa_run_static: the same result as jfa_run, computed in the library's static device scratch (no device
    allocation in the call; images of at most JFA_STATIC_MAX_PIXELS pixels; calls on one GPU must not overlap)."""
    seeds = _img(seeds, np.uint32, 2)
    h, w, _ = seeds.shape
    out = np.empty_like(seeds)
    _lib.check(_lib.load().fluid_jfa_run_static(_lib.ptr(seeds), w, h, _lib.ptr(out), device, None, 1))
    return out


def jfa_pass(seeds: np.ndarray, jump: int, device: int = 0) -> np.ndarray:
    seeds = _img(seeds, np.uint32, 2)
    h, w, _ = seeds.shape
    out = np.empty_like(seeds)
    _lib.check(_lib.load().fluid_jfa_pass(_lib.ptr(seeds), w, h, int(jump), _lib.ptr(out), device))
    return out


def seed_from_mask(mask: np.ndarray, device: int = 0) -> np.ndarray:
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    h, w = mask.shape
    out = np.empty((h, w, 2), dtype=np.uint32)
    _lib.check(_lib.load().fluid_jfa_seed_from_mask(_lib.ptr(mask), w, h, _lib.ptr(out), device))
    return out


def seed_from_particles(particles: np.ndarray, bounds, w: int, h: int, device: int = 0) -> np.ndarray:
    particles = np.ascontiguousarray(particles)
    assert particles.dtype == _lib.PARTICLE_DTYPE
    b = (C.c_float * 2)(float(bounds[0]), float(bounds[1]))
    out = np.empty((h, w, 2), dtype=np.uint32)
    _lib.check(_lib.load().fluid_jfa_seed_from_particles(_lib.ptr(particles), particles.shape[0], C.byref(b), w, h,
                                                         _lib.ptr(out), device))
    return out


def to_force_field(seeds: np.ndarray, device: int = 0) -> np.ndarray:
    seeds = _img(seeds, np.uint32, 2)
    h, w, _ = seeds.shape
    out = np.empty((h, w, 2), dtype=np.float32)
    _lib.check(_lib.load().fluid_jfa_to_force_field(_lib.ptr(seeds), w, h, _lib.ptr(out), device))
    return out

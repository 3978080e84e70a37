"""Obstacle objects -> mask -> seeds -> JFA -> force field: host mirror of the reference's
object pipeline (src/renderer.rs:60-90, :449-547, image_shader.wgsl). All compute runs in libfluid_b200.so."""
import ctypes as C

import numpy as np

from . import _lib

OBJECT_DTYPE = _lib.OBJECT_DTYPE


def _bits(x) -> int:
    return int(np.array([x], dtype=np.float32).view(np.uint32)[0])


def circle(position, radius):
    """FluidObject kind 0 (image_shader.wgsl:50-55): radius travels bit-cast in `pad`."""
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["position"] = position
    o["kind"] = 0
    o["pad"] = _bits(radius)
    return o


def rect(position, extents, rotation=0.0):
    """FluidObject kind 1 (image_shader.wgsl:56-61): rotation in `pad`, extents bit-cast in pad2.xy."""
    o = np.zeros((), dtype=OBJECT_DTYPE)
    o["position"] = position
    o["kind"] = 1
    o["pad"] = _bits(rotation)
    o["pad2"] = (_bits(extents[0]), _bits(extents[1]), 0, 0)
    return o


def objects(*objs) -> np.ndarray:
    return np.array(list(objs), dtype=OBJECT_DTYPE) if objs else np.zeros(0, dtype=OBJECT_DTYPE)


def inverse_projection(size) -> np.ndarray:
    """inverse of glam::Mat4::orthographic_rh(-sx/2, sx/2, sy/2, -sy/2, -1, 0) (src/renderer.rs:558-561),
    16 floats, column-major — what the reference puts in ObjectUniform.inv_proj (:449-453)."""
    sx, sy = float(size[0]), float(size[1])
    left, right, bottom, top, near, far = -sx * 0.5, sx * 0.5, sy * 0.5, -sy * 0.5, -1.0, 0.0
    rw, rh, r = 1.0 / (right - left), 1.0 / (top - bottom), 1.0 / (near - far)
    m = np.array([[2 * rw, 0, 0, -(left + right) * rw],
                  [0, 2 * rh, 0, -(top + bottom) * rh],
                  [0, 0, r, r * near],
                  [0, 0, 0, 1]], dtype=np.float64)  # row-major view of the column vectors
    return np.ascontiguousarray(np.linalg.inv(m).T.astype(np.float32)).reshape(16)  # flatten column-major


def raster_objects(objs: np.ndarray, inv_proj, w: int, h: int, device: int = 0) -> np.ndarray:
    """mask only, host buffers: u8[h, w], 0 inside an object, 255 elsewhere."""
    lib = _lib.load()
    objs = np.ascontiguousarray(objs, dtype=OBJECT_DTYPE)
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    out = np.empty((h, w), dtype=np.uint8)
    _lib.check(lib.fluid_raster_objects(_lib.ptr(objs), objs.shape[0], _lib.ptr(m), w, h, _lib.ptr(out), device))
    return out


def field_from_objects(sim, objs: np.ndarray, inv_proj, want_mask=False, want_seeds=False, want_field=False, timed=False):
    """device-resident chain on `sim` (a FluidSimulation): the field is installed as the sim's force field.
    Returns (mask, seeds, field, ms) with None for outputs not requested; with nothing requested the call only
    enqueues work on the sim's stream."""
    tw, th = sim.settings.texture_size
    objs = np.ascontiguousarray(objs, dtype=OBJECT_DTYPE)
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    mask = np.empty((th, tw), dtype=np.uint8) if want_mask else None
    seeds = np.empty((th, tw, 2), dtype=np.uint32) if want_seeds else None
    field = np.empty((th, tw, 2), dtype=np.float32) if want_field else None
    ms = C.c_float(0.0)
    _lib.check(sim._lib.fluid_sim_field_from_objects(sim._h, _lib.ptr(objs), objs.shape[0], _lib.ptr(m), _lib.ptr(mask),
                                                    _lib.ptr(seeds), _lib.ptr(field), C.byref(ms) if timed else None))
    return mask, seeds, field, (ms.value if timed else None)


def field_from_mask(sim, mask: np.ndarray, want_seeds=False, want_field=False, timed=False):
    """the same from a caller-supplied u8 mask at the sim's texture size (the reference's `Image`)."""
    tw, th = sim.settings.texture_size
    mask = np.ascontiguousarray(mask, dtype=np.uint8)
    if mask.shape != (th, tw):
        raise ValueError(f"mask must be {(th, tw)}, got {mask.shape}")
    seeds = np.empty((th, tw, 2), dtype=np.uint32) if want_seeds else None
    field = np.empty((th, tw, 2), dtype=np.float32) if want_field else None
    ms = C.c_float(0.0)
    _lib.check(sim._lib.fluid_sim_field_from_mask(sim._h, _lib.ptr(mask), _lib.ptr(seeds), _lib.ptr(field),
                                                 C.byref(ms) if timed else None))
    return seeds, field, (ms.value if timed else None)


def render(sim, inv_proj, w: int = 960, h: int = 540, want_f32=True, want_rgba8=False, timed=False):
    """headless fluid_shader.wgsl: returns (float32[h, w, 4] fragment output, uint8[h, w, 4] blended image, ms)."""
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    f = np.empty((h, w, 4), dtype=np.float32) if want_f32 else None
    b = np.empty((h, w, 4), dtype=np.uint8) if want_rgba8 else None
    ms = C.c_float(0.0)
    _lib.check(sim._lib.fluid_sim_render(sim._h, _lib.ptr(m), int(w), int(h), _lib.ptr(f), _lib.ptr(b),
                                        C.byref(ms) if timed else None))
    return f, b, (ms.value if timed else None)


def surface_tension(sim) -> np.ndarray:
    """row f4: calculate_surface_tension (compute.wgsl:300-498) for every sorted particle of the last tick's state;
    float32[n, 2]. The reference's call is commented out (compute.wgsl:92), so nothing is applied to the sim."""
    out = np.empty((sim.particle_count, 2), dtype=np.float32)
    _lib.check(sim._lib.fluid_sim_surface_tension(sim._h, _lib.ptr(out)))
    return out


def mouse_world_pos(cursor_xy, window_size, inv_proj) -> np.ndarray:
    """CursorMoved mapping of src/main.rs:287-303: window pixel -> TickSettings.mouse_pos (world space)."""
    c = np.ascontiguousarray(cursor_xy, dtype=np.float32).reshape(2)
    w = np.ascontiguousarray(window_size, dtype=np.float32).reshape(2)
    m = np.ascontiguousarray(inv_proj, dtype=np.float32).reshape(16)
    out = np.empty(2, dtype=np.float32)
    _lib.check(_lib.load().fluid_mouse_world_pos(_lib.ptr(c), _lib.ptr(w), _lib.ptr(m), _lib.ptr(out)))
    return out

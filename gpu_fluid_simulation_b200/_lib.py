"""ctypes binding of libfluid_b200.so (the C ABI of include/fluid_b200.h).

There is no CPU fallback: if the shared library is missing or does not load, importing the
product path raises. PyTorch is not involved at this level.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLUID_B200_LIB") or os.path.join(HERE, "libfluid_b200.so")  # (the override is for A/B runs of experimental builds)

# numpy twins of the C structs (reference layouts: src/simulation.rs:53-135)
PARTICLE_DTYPE = np.dtype([
    ("position", np.float32, 2), ("predicted_position", np.float32, 2), ("velocity", np.float32, 2),
    ("density", np.float32), ("grid", np.uint32)], align=False)
assert PARTICLE_DTYPE.itemsize == 32
# reference FluidObject (src/renderer.rs:82-90): pad / pad2 carry bit-cast floats
OBJECT_DTYPE = np.dtype([("position", np.float32, 2), ("kind", np.uint32), ("pad", np.uint32), ("pad2", np.uint32, 4)], align=False)
assert OBJECT_DTYPE.itemsize == 32


class SimSettings(C.Structure):
    _fields_ = [("particle_count", C.c_uint32), ("particle_spacing", C.c_float), ("smoothing_radius", C.c_float),
                ("size", C.c_float * 2), ("texture_size", C.c_uint32 * 2)]


class TickSettingsC(C.Structure):
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


assert C.sizeof(Uniform) == 120 and C.sizeof(SimSettings) == 28 and C.sizeof(TickSettingsC) == 60


class DeviceView(C.Structure):
    _fields_ = [("particle_count", C.c_uint32), ("capacity", C.c_uint32), ("position", C.c_void_p),
                ("predicted_position", C.c_void_p), ("velocity", C.c_void_p), ("density", C.c_void_p),
                ("grid", C.c_void_p), ("start_indices", C.c_void_p), ("cell_lower_bound", C.c_void_p),
                ("force_field", C.c_void_p)]


class MultiOptions(C.Structure):
    """FluidMultiOptions (include/fluid_b200.h)."""
    _fields_ = [("row_bounds", C.c_uint32 * 9), ("capacity", C.c_uint32), ("window_near", C.c_uint32),
                ("window_far", C.c_uint32), ("transport", C.c_int32), ("reserved", C.c_uint32 * 4)]


_P = C.c_void_p
_u32 = C.c_uint32
_i = C.c_int

# every symbol include/fluid_b200.h declares: name -> (restype, argtypes)
PROTOTYPES = {
    "fluid_last_error": (C.c_char_p, []),
    "fluid_abi_version": (_i, []),
    "fluid_device_count": (_i, []),
    "fluid_sim_create": (_i, [C.POINTER(SimSettings), _i, C.POINTER(_P)]),
    "fluid_sim_destroy": (None, [_P]),
    "fluid_sim_tick": (_i, [_P, C.POINTER(TickSettingsC)]),
    "fluid_sim_tick_n": (_i, [_P, C.POINTER(TickSettingsC), _u32]),
    "fluid_sim_sync": (_i, [_P]),
    "fluid_sim_tick_count": (_u32, [_P]),
    "fluid_sim_set_tick_count": (_i, [_P, _u32]),
    "fluid_sim_set_math_mode": (_i, [_P, _i]),
    "fluid_sim_get_math_mode": (_i, [_P]),
    "fluid_sim_set_sort_mode": (_i, [_P, _i]),
    "fluid_sim_get_sort_mode": (_i, [_P]),
    "fluid_sim_last_sort_path": (_i, [_P]),
    "fluid_sim_tick_hash_sort_only": (_i, [_P, C.POINTER(TickSettingsC)]),
    "fluid_sim_particle_count": (_u32, [_P]),
    "fluid_sim_grid_w": (_u32, [_P]),
    "fluid_sim_grid_h": (_u32, [_P]),
    "fluid_sim_download_particles": (_i, [_P, _P]),
    "fluid_sim_upload_particles": (_i, [_P, _P]),
    "fluid_sim_upload_particles_async": (_i, [_P, _P]),
    "fluid_sim_download_particles_async": (_i, [_P, _P]),
    "fluid_sim_sync_transfers": (_i, [_P]),
    "fluid_sim_download_start_indices": (_i, [_P, _P]),
    "fluid_sim_upload_start_indices": (_i, [_P, _P]),
    "fluid_sim_upload_force_field": (_i, [_P, _P]),
    "fluid_sim_download_force_field": (_i, [_P, _P]),
    "fluid_sim_get_uniform": (_i, [_P, C.POINTER(Uniform)]),
    "fluid_sim_download_unsorted_keys": (_i, [_P, _P]),
    "fluid_sim_enable_debug_force": (_i, [_P, _i]),
    "fluid_sim_download_debug_force": (_i, [_P, _P, _P]),
    "fluid_sim_download_permutation": (_i, [_P, _P]),
    "fluid_sim_device_view": (_i, [_P, C.POINTER(DeviceView)]),
    "fluid_sim_state_changed": (_i, [_P]),
    "fluid_sim_stream": (_P, [_P]),
    "fluid_sim_enable_timing": (_i, [_P, _i]),
    "fluid_sim_last_timing": (_i, [_P, C.POINTER(C.c_char_p), C.POINTER(C.c_float), _i]),
    "fluid_sim_launch_count": (C.c_uint64, [_P]),
    "fluid_sim_reserve": (_i, [_P, _u32]),
    "fluid_sim_set_particle_count": (_i, [_P, _u32]),
    "fluid_multi_create": (_i, [C.POINTER(SimSettings), _i, C.POINTER(_i), C.POINTER(MultiOptions), C.POINTER(_P)]),
    "fluid_multi_unique_id": (_i, [_P]),
    "fluid_multi_create_rank": (_i, [C.POINTER(SimSettings), _i, _i, _P, _i, C.POINTER(MultiOptions), C.POINTER(_P)]),
    "fluid_multi_destroy": (None, [_P]),
    "fluid_multi_tick": (_i, [_P, C.POINTER(TickSettingsC)]),
    "fluid_multi_tick_n": (_i, [_P, C.POINTER(TickSettingsC), _u32]),
    "fluid_multi_sync": (_i, [_P]),
    "fluid_multi_tick_count": (_u32, [_P]),
    "fluid_multi_set_tick_count": (_i, [_P, _u32]),
    "fluid_multi_set_math_mode": (_i, [_P, _i]),
    "fluid_multi_world": (_i, [_P]),
    "fluid_multi_local_bands": (_i, [_P]),
    "fluid_multi_transport": (_i, [_P]),
    "fluid_multi_row_bounds": (_i, [_P, _P]),
    "fluid_multi_upload_particles": (_i, [_P, _P, _u32]),
    "fluid_multi_download_particles": (_i, [_P, _P]),
    "fluid_multi_upload_band": (_i, [_P, _P, _u32, _P]),
    "fluid_multi_download_band": (_i, [_P, _i, _P, _u32, C.POINTER(_u32), C.POINTER(_u32)]),
    "fluid_multi_upload_band_async": (_i, [_P, _P, _u32]),
    "fluid_multi_download_band_async": (_i, [_P, _P, _u32, _P]),
    "fluid_multi_sync_transfers": (_i, [_P]),
    "fluid_multi_particle_counts": (_i, [_P, _P]),
    "fluid_multi_download_start_indices": (_i, [_P, _P]),
    "fluid_multi_upload_start_indices": (_i, [_P, _P]),
    "fluid_multi_upload_force_field": (_i, [_P, _P]),
    "fluid_multi_enable_timing": (_i, [_P, _i]),
    "fluid_multi_last_timing": (_i, [_P, _i, C.POINTER(C.c_char_p), C.POINTER(C.c_float), _i]),
    "fluid_multi_launch_count": (C.c_uint64, [_P]),
    "fluid_multi_stream": (_P, [_P, _i]),
    "fluid_multi_band_sim": (_P, [_P, _i]),
    "fluid_multi_balanced_bounds": (_i, [C.POINTER(SimSettings), _P, _u32, _i, _P]),
    "fluid_jfa_run": (_i, [_P, _u32, _u32, _P, _i]),
    "fluid_jfa_pass": (_i, [_P, _u32, _u32, _u32, _P, _i]),
    "fluid_jfa_run_static": (_i, [_P, _u32, _u32, _P, _i, _P, _i]),
    "fluid_jfa_seed_from_mask": (_i, [_P, _u32, _u32, _P, _i]),
    "fluid_jfa_seed_from_particles": (_i, [_P, _u32, C.POINTER(C.c_float * 2), _u32, _u32, _P, _i]),
    "fluid_jfa_to_force_field": (_i, [_P, _u32, _u32, _P, _i]),
    "fluid_sim_jfa_from_particles": (_i, [_P, _u32, _u32, _P, _P, _i, C.POINTER(C.c_float)]),
    "fluid_raster_objects": (_i, [_P, _u32, _P, _u32, _u32, _P, _i]),
    "fluid_sim_field_from_objects": (_i, [_P, _P, _u32, _P, _P, _P, _P, C.POINTER(C.c_float)]),
    "fluid_sim_field_from_mask": (_i, [_P, _P, _P, _P, C.POINTER(C.c_float)]),
    "fluid_sim_render": (_i, [_P, _P, _u32, _u32, _P, _P, C.POINTER(C.c_float)]),
    "fluid_sim_render_frame": (_i, [_P, _P, _u32, _u32, _P, _i, _i]),
    "fluid_sim_surface_tension": (_i, [_P, _P]),
    "fluid_mouse_world_pos": (_i, [_P, _P, _P, _P]),
    "fluid_selftest_division": (_i, [_i, _u32, _i, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                     C.POINTER(C.c_float)]),
}

_lib = None


class FluidError(RuntimeError):
    pass


def load():
    """Load libfluid_b200.so; raises if it is missing (no fallback of any kind)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FluidError(
            f"{LIB_PATH} is missing. Build it with `python -m gpu_fluid_simulation_b200._build` "
            "(needs nvcc). There is no CPU fallback for the fluid step.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().fluid_last_error()
        raise FluidError(f"libfluid_b200 status {rc}: {msg.decode() if msg else ''}")


def ptr(arr):
    """void* of a C-contiguous numpy array, or of a torch tensor / raw address."""
    if arr is None:
        return None
    if isinstance(arr, int):
        return C.c_void_p(arr)
    if isinstance(arr, np.ndarray):
        if not arr.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return C.c_void_p(arr.ctypes.data)
    if hasattr(arr, "data_ptr"):
        return C.c_void_p(arr.data_ptr())
    raise TypeError(type(arr))

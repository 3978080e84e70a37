"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol the header
declares, has the reference's struct layouts, fails loudly without a GPU, and never touches the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "fluid_b200.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(fluid_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_symbols_all_exported(fluid):
    from gpu_fluid_simulation_b200 import _lib
    lib = fluid.load()
    syms = header_symbols()
    assert len(syms) >= 35
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r" T (fluid_[a-z0-9_]+)", out))
    missing = [s for s in syms if s not in exported]
    assert not missing, f"declared in include/fluid_b200.h but not exported: {missing}"
    # and the Python prototypes cover the same set
    assert sorted(_lib.PROTOTYPES) == syms
    assert lib.fluid_abi_version() == 2


def test_struct_layouts_match_reference(fluid):
    from gpu_fluid_simulation_b200 import _lib
    assert _lib.PARTICLE_DTYPE.itemsize == 32
    assert [_lib.PARTICLE_DTYPE.fields[k][1] for k in ("position", "predicted_position", "velocity", "density", "grid")] == [0, 8, 16, 24, 28]
    assert C.sizeof(_lib.Uniform) == 120
    U = _lib.Uniform
    offs = dict(delta=0, particle_count=4, sqr_radius=8, frame_time=12, gravity=16, bounds=24, mouse_pos=32,
                smoothing_radius=40, particle_mass=44, pressure_constant=48, rest_density=52, damping_factor=56,
                viscosity_coefficient=60, surface_tension_treshold=64, surface_tension_coefficient=68,
                poly6_kernel_volume=72, poly6_kernel_derivative=76, poly6_kernel_laplacian=80,
                spiky_kernel_derivative=84, viscosity_kernel=88, mouse_state=92, mouse_force_radius=96,
                mouse_force_power=100, grid_w=104, grid_h=108, texture_size=112)
    for k, v in offs.items():
        assert getattr(U, k).offset == v, k


def test_no_gpu_fails_loudly(fluid):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(fluid.FluidError) as e:
        fluid.FluidSimulation.new(0, fluid.SimulationSettings())
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(fluid.FluidError):
        fluid.jfa.jfa_run(np.zeros((4, 4, 2), np.uint32))


def test_argument_validation_precedes_device_use(fluid):
    lib = fluid.load()
    from gpu_fluid_simulation_b200 import _lib
    h = C.c_void_p()
    s = fluid.SimulationSettings(particle_count=1).to_c()
    assert lib.fluid_sim_create(C.byref(s), 0, C.byref(h)) == -1  # the reference panics for N <= 1 (src/simulation.rs:323-324)
    assert b"particle_count < 2" in lib.fluid_last_error()
    assert lib.fluid_sim_create(None, 0, C.byref(h)) == -1
    assert lib.fluid_sim_tick(None, None) == -1
    assert lib.fluid_sim_tick_count(None) == 0


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gpu_fluid_simulation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inl", ".c", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.lower(), f"{f} mentions the oracle"


def test_scenes_deterministic(fluid):
    from gpu_fluid_simulation_b200 import scenes
    s, ts, p = scenes.uniform_box(32)
    s2, _, p2 = scenes.uniform_box(32)
    assert np.array_equal(p, p2) and s == s2
    scenes.assert_no_coincident(p)
    assert abs(float(np.abs(p["velocity"]).max()) - 1.0) < 0.01
    s, ts, p = scenes.dam_break(16, 64)
    assert ts.gravity == (0.0, 9.8) and s.size == (16 * 0.1 * 4.0, 64 * 0.1)
    assert p["position"][:, 0].max() < -s.size[0] / 4 + 0.05
    r = scenes.splitmix64(1, 4)
    assert (r >= 0).all() and (r < 1).all() and len(set(r.tolist())) == 4


def test_header_constants_match_the_python_mirror(fluid):
    """the #defines a C / Rust host reads are the numbers the Python mirror uses (transports, limits, ABI version)."""
    import re
    from gpu_fluid_simulation_b200 import jfa, multi
    with open(os.path.join(ROOT, "include", "fluid_b200.h")) as f:
        text = f.read()
    defs = {m.group(1): m.group(2) for m in re.finditer(r"^#define\s+(FLUID_\w+)\s+\(?([0-9a-fxu<> ]+?)\)?\s*(?:/\*.*)?$", text, flags=re.M)}

    def val(name):
        return int(eval(defs[name].replace("u", "")))  # "1u << 20" -> 1 << 20
    assert val("FLUID_TRANSPORT_AUTO") == multi.TRANSPORT_AUTO
    assert val("FLUID_TRANSPORT_NCCL") == multi.TRANSPORT_NCCL
    assert val("FLUID_TRANSPORT_LOCAL") == multi.TRANSPORT_LOCAL
    assert val("FLUID_TRANSPORT_PEER") == multi.TRANSPORT_PEER
    assert val("FLUID_JFA_STATIC_MAX_PIXELS") == jfa.JFA_STATIC_MAX_PIXELS
    assert val("FLUID_MULTI_MAX_RANKS") == 8

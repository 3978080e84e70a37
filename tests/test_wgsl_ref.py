"""Pins the hand-written oracle (oracle/fluid_oracle.cpp) to the reference by a second, MECHANICAL derivation: the
reference's own WGSL (compute.wgsl + funcs.wgsl, sort.wgsl, shaders/jump_flood.wgsl), rewritten token by token into
C++ by oracle/transpile_wgsl.py and compiled into oracle/_ref/libwgsl_ref.so. Every comparison is bit for bit.

The reference has no tests or golden vectors and cannot run here (Rust + wgpu + a graphics API), so this is the closest
thing to "the reference compiled here" the box allows. What it does NOT cover: the Rust host code (uniform derivation,
init lattice, bitonic schedule, generate_smooth_gradient_field) — those stay restated by hand in both libraries."""
import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
from oracle import wgsl_ref as W  # noqa: E402

pytestmark = pytest.mark.skipif(not W.available(), reason="oracle/_ref/libwgsl_ref.so not built and /root/reference absent")


def _golden_cases():
    import make_golden
    for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "*.npz"))):
        z = np.load(path)
        if "settings" not in z.files:
            continue
        s, ts = make_golden.load_settings(z)
        yield os.path.basename(path), s, ts, z


def _same(a, b, what):
    assert a.tobytes() == b.tobytes(), f"{what}: transpiled WGSL and hand-written oracle differ"


@pytest.mark.parametrize("name", [c[0] for c in _golden_cases()])
def test_full_ticks_on_the_golden_scenes(oracle, name):
    """whole ticks (predict, key, the 100+-pass bitonic network, cell table, 7x7 density, move) from each golden input:
    every byte of every particle and of start_indices."""
    O = oracle
    _, s, ts, z = next(c for c in _golden_cases() if c[0] == name)
    a = O.OracleSim(s, z["input"], sort_mode=O.SORT_BITONIC, walk_radius=3)
    b = W.WgslRefSim(O, s, z["input"])
    if "field" in z.files:
        a.texture[:] = z["field"]; b.texture[:] = z["field"]
    if "table_in" in z.files:
        a.table[:] = z["table_in"]; b.table[:] = z["table_in"]
    for k in range(int(z["ticks"][0])):
        a.tick(ts)
        b.tick(ts)
        _same(a.p, b.p, f"{name} tick {k + 1} particles")
        _same(a.table, b.table, f"{name} tick {k + 1} start_indices")
    # and the algorithm-independent parts equal the committed golden output (made with the STABLE sort): sorted keys and
    # the cell table. Not for the sheared default lattice: its coincident pairs take the xorshift branch seeded with the
    # sorted index (compute.wgsl:161), so the bitonic network's tie order changes their positions from tick 2 on.
    pos = np.ascontiguousarray(z["input"]["position"]).view(np.uint64).ravel()
    if np.unique(pos).size == pos.size:
        assert np.array_equal(b.p["grid"], z["output"]["grid"])
        assert np.array_equal(b.table, z["table"])


@pytest.mark.parametrize("n", [2, 3, 1000, 4097])
def test_bitonic_network(oracle, n):
    """sort.wgsl under the schedule of src/simulation.rs:323-347: the whole records, i.e. also the (unstable) tie order."""
    rng = np.random.default_rng(n)
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["grid"] = rng.integers(0, max(2, n // 3), size=n)
    p["density"] = np.arange(n, dtype=np.float32)  # tags the records
    a, b = p.copy(), p.copy()
    oracle.sort_bitonic(a)
    W.sort(b)
    _same(a, b, "bitonic network")
    assert np.array_equal(b["grid"], np.sort(p["grid"]))


def test_single_passes_with_mouse_field_walls_and_coincident_particles(oracle):
    O = oracle
    from gpu_fluid_simulation_b200 import scenes
    s, ts, p = scenes.uniform_box(36, tex=(16, 16))
    ts.mouse_state, ts.mouse_pos, ts.mouse_force_radius = 1, (0.4, 0.1), 1.0
    ts.gravity = (1.5, 9.8)
    rng = np.random.default_rng(2)
    p["velocity"] *= 40.0                                  # wall hits within a tick
    p[5]["position"] = p[9]["position"]                    # a coincident pair -> the xorshift branch (compute.wgsl:211-212)
    p[5]["velocity"] = p[9]["velocity"]
    field = rng.integers(-1, 2, size=(16, 16, 2)).astype(np.float32)
    gw, gh = O.grid_dims(s)
    table = rng.integers(0, p.shape[0], size=gw * gh).astype(np.uint32)  # stale garbage (never-cleared table)
    u = O.make_uniform(s, ts, 7)
    a, b = p.copy(), p.copy()
    ta, tb = table.copy(), table.copy()
    O.predict_next_position(u, a); W.run_pass("predict_next_position", u, b, tb, field)
    _same(a, b, "predict_next_position")
    O.create_spatial_lookup(u, a); W.run_pass("create_spatial_lookup", u, b, tb, field)
    _same(a, b, "create_spatial_lookup")
    O.sort_bitonic(a); W.sort(b)
    _same(a, b, "sort")
    O.compute_start_indices(u, a, ta); W.run_pass("compute_start_indices", u, b, tb, field)
    _same(ta, tb, "compute_start_indices")
    O.calculate_density(u, a, ta, 3); W.run_pass("calculate_density", u, b, tb, field)
    _same(a, b, "calculate_density")
    a2 = O.move_particle(u, a, ta, field)
    b2 = W.move_particle(u, b, tb, field)
    _same(a2, b2, "move_particle")
    assert (np.abs(a2["position"][:, 0]) == np.float32(s.size[0] / 2)).any(), "no particle reached a wall: weak case"


def test_jump_flood_pass(oracle):
    rng = np.random.default_rng(4)
    for (h, w) in ((64, 64), (40, 96), (96, 40)):
        seeds = np.full((h, w, 2), 32767, np.uint32)
        ys, xs = rng.integers(0, h, 30), rng.integers(0, w, 30)
        seeds[ys, xs, 0], seeds[ys, xs, 1] = xs, ys
        seeds[0, 0] = (w + 5, 3)  # a value jump_flood.wgsl:48 rejects as a neighbour but keeps as an incumbent
        cur_a, cur_b = seeds.copy(), seeds.copy()
        j = 1
        while j * 2 < max(w, h):
            j *= 2
        while j >= 1:
            cur_a = oracle.jfa_pass(cur_a, j)
            cur_b = W.jfa_pass(cur_b, j)
            _same(cur_a, cur_b, f"jfa pass {w}x{h} jump {j}")
            j //= 2
        _same(oracle.jfa_run(seeds), cur_b, "jfa full run")

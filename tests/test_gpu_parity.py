"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle from identical input states.

Criteria (BASELINE.json north star): (1) hash keys, sorted keys, cell tables bit-exact; (2) density, velocity,
position within fp32 relative 1e-5 (force-like quantities relative to sum|terms|) — and, because the radix sort's
tie order equals the oracle's stable order and every fp32 operation is issued in the reference's order without
contraction, actually bit-identical, which is asserted where noted."""
import os

import numpy as np
import pytest

import parity
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _check(fluid, oracle, s, ts, p, **kw):
    osim, sim, extras = parity.run_pair(fluid, oracle, s, ts, p, **kw)
    try:
        stats = parity.compare_state(osim, sim, extras, ts)
    finally:
        sim.close()
    return stats


@pytest.mark.parametrize("n_side,ticks", [(16, 1), (64, 1), (200, 1), (128, 8)])
def test_uniform_box(fluid, oracle, n_side, ticks):
    s, ts, p = scenes.uniform_box(n_side)
    scenes.assert_no_coincident(p)
    stats = _check(fluid, oracle, s, ts, p, n_ticks=ticks)
    assert stats["floats_bit_exact"], stats


@pytest.mark.parametrize("cols,rows,ticks", [(32, 128, 1), (64, 256, 30)])
def test_dam_break(fluid, oracle, cols, rows, ticks):
    s, ts, p = scenes.dam_break(cols, rows)
    stats = _check(fluid, oracle, s, ts, p, n_ticks=ticks)
    assert stats["floats_bit_exact"], stats


def test_default_scene_with_coincident_pairs(fluid, oracle):
    """reference default scene (src/main.rs:48-54): 72 coincident pairs exercise the xorshift branch, whose seed is
    the sorted index — identical here because the tie order is the oracle's stable order."""
    s = SimulationSettings()
    p = oracle.init_lattice(s)
    stats = _check(fluid, oracle, s, TickSettings(), p, n_ticks=3)
    assert stats["floats_bit_exact"], stats


@pytest.mark.parametrize("scene", ["U", "D", "default", "field_mouse"])
def test_fast_math_mode_within_tolerance(fluid, oracle, scene):
    """FLUID_MATH_FAST: hardware rsqrt/rcp + FMA in the pair terms. Keys, sorted keys, tables, predicted positions and
    densities stay bit-exact; acceleration within 1e-5 of sum|terms|; velocity / position within 1e-5 (criterion 2)."""
    field = None
    if scene == "U":
        s, ts, p = scenes.uniform_box(200)
    elif scene == "D":
        s, ts, p = scenes.dam_break(64, 256)
    elif scene == "default":
        s = SimulationSettings(); ts = TickSettings(); p = oracle.init_lattice(s)
    else:
        s, ts, p = scenes.uniform_box(64, tex=(32, 32))
        rng = np.random.default_rng(9)
        field = rng.integers(-4, 5, size=(32, 32, 2)).astype(np.float32)
        ts.mouse_state = 1; ts.mouse_pos = (0.7, -1.1); ts.mouse_force_radius = 2.0
    osim, sim, extras = parity.run_pair(fluid, oracle, s, ts, p, field=field, n_ticks=1, math_mode=1)
    try:
        assert sim.math_mode == 1
        stats = parity.compare_state(osim, sim, extras, ts)
        got = sim.download_particles()
    finally:
        sim.close()
    assert np.array_equal(got["density"], osim.p["density"])
    assert not stats["floats_bit_exact"] or scene == "x"  # it really is a different arithmetic path
    assert stats["accel_max_rel_to_terms"] <= 1e-5 and stats["velocity_max_rel"] <= 1e-5, stats


def test_create_builds_reference_lattice(fluid, oracle):
    s = SimulationSettings(particle_count=12345, size=(20.0, 20.0))
    with fluid.FluidSimulation.new(0, s) as sim:
        assert np.array_equal(sim.download_particles(), oracle.init_lattice(s))
        assert (sim.grid_w, sim.grid_h) == oracle.grid_dims(s)
        assert sim.tick_count == 0
        assert not sim.download_start_indices().any()
        sim.tick(TickSettings())
        assert sim.tick_count == 1 and sim.uniform().frame_time == 1


@pytest.mark.parametrize("n", [2, 3, 255, 256, 257, 4097, 8191])
def test_ragged_counts(fluid, oracle, n):
    rng = np.random.default_rng(n)
    s = SimulationSettings(particle_count=n, size=(6.0, 4.0), texture_size=(16, 16))
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = (rng.random((n, 2)) - 0.5) * np.array([6.0, 4.0]) * 0.98
    p["velocity"] = rng.normal(size=(n, 2)) * 3
    p["predicted_position"] = p["position"]
    stats = _check(fluid, oracle, s, TickSettings(gravity=(0.0, 9.8)), p, n_ticks=2)
    assert stats["floats_bit_exact"], stats


def test_all_particles_in_one_cell_and_exact_duplicates(fluid, oracle):
    n = 700
    s = SimulationSettings(particle_count=n, size=(4.0, 4.0), texture_size=(8, 8))
    rng = np.random.default_rng(1)
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = 0.05 + rng.random((n, 2)) * 0.1
    p["position"][100:140] = p["position"][0]  # 41 exactly coincident particles -> RNG directions
    p["predicted_position"] = p["position"]
    stats = _check(fluid, oracle, s, TickSettings(), p, n_ticks=2, tick_count=41)
    assert stats["floats_bit_exact"], stats


@pytest.mark.parametrize("n,extent,sort_mode", [(6000, (0.9, 0.5), 0), (6000, (0.9, 0.5), 1), (40000, (0.5, 0.15), 0)])
def test_neighbourhoods_larger_than_the_staging_buffer(fluid, oracle, n, extent, sort_mode):
    """the late dam break in small: cells of hundreds of particles, so the three row spans of a 256-particle tile exceed the
    staging buffer (kernels_sph.cuh SPH_CAP) and the tile walks global memory — with the accepted-candidate list while
    every span is shorter than 2^14, with the plain walk beyond (the 40000-particle case: ~13000 per cell). The
    reference walks any population (compute.wgsl:173-232). Includes exactly coincident particles (RNG directions)."""
    s = SimulationSettings(particle_count=n, size=(4.0, 4.0), texture_size=(8, 8))
    rng = np.random.default_rng(n)
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = np.float32(0.31) + rng.random((n, 2)) * np.array(extent)  # straddles cell boundaries (h = 0.2)
    p["position"][200:215] = p["position"][3]
    p["velocity"] = rng.normal(size=(n, 2)) * 0.5
    p["predicted_position"] = p["position"]
    stats = _check(fluid, oracle, s, TickSettings(), p, n_ticks=2, tick_count=7, gpu_sort_mode=sort_mode)
    assert stats["floats_bit_exact"], stats


@pytest.mark.parametrize("wild", ["tiny_velocity", "tiny_offset", "huge_mass", "huge_stiffness", "none"])
def test_range_tested_path_for_wild_inputs(fluid, oracle, wild):
    """the force pass skips most per-pair range tests of its shared-reciprocal divisions when the tick's inputs are
    "tame" (components zero or in [2^-70, 2^99], densities <= 2^20: stamped by the gather and density passes). One wild
    value anywhere must send the whole tick through the fully range-tested body — same bits either way."""
    s, ts, p = scenes.uniform_box(64)
    if wild == "tiny_velocity":
        p["velocity"][1234] = [3e-30, -7e-33]          # differences of ~1e-30: below the 2^-100 numerator bound
        p["velocity"][1235] = [0.0, 1e-38]
    elif wild == "tiny_offset":
        # two particles 1e-30 apart in x next to the origin: dx tiny but non-zero, r < 2^-20
        p["position"][2000] = [1e-30, 2e-31]
        p["position"][2001] = [2e-30, 2e-31]
        p["predicted_position"][2000:2002] = p["position"][2000:2002]
        p["velocity"][2000:2002] = 0
    elif wild == "huge_mass":
        ts = TickSettings(**{**ts.__dict__, "mass": 3e8})  # densities above 2^20
    elif wild == "huge_stiffness":
        ts = TickSettings(**{**ts.__dict__, "pressure_constant": 1e21})  # the per-launch bound on the pressure numerators fails
    stats = _check(fluid, oracle, s, ts, p, n_ticks=2)
    assert stats["floats_bit_exact"], stats


def test_walls_speed_clamp_and_nan_guard(fluid, oracle):
    n = 2048
    s = SimulationSettings(particle_count=n, size=(8.0, 8.0), texture_size=(8, 8))
    rng = np.random.default_rng(3)
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = (rng.random((n, 2)) - 0.5) * 8.0
    p["velocity"] = rng.normal(size=(n, 2)) * 400  # many exceed 500 and cross the walls in one step
    # (NaN / inf *inputs* are outside the contract: WGSL leaves NaN behaviour implementation-defined, an inf velocity
    # turns into a NaN position through the speed clamp (inf/inf) in the reference too, and the 7x7 density walk then
    # sees that NaN from further away than a 3x3 walk does. The NaN guard itself (compute.wgsl:113-116) is exercised by
    # the mouse test, where dist == 0 produces NaN inside the tick.)
    p["position"][7] = [4.0, -4.0]  # exactly on the corner: uv == 1.0 texel overflow (compute.wgsl:127-129)
    p["predicted_position"] = p["position"]
    osim, sim, extras = parity.run_pair(fluid, oracle, s, TickSettings(gravity=(0.0, 9.8)), p, n_ticks=2, debug_force=False)
    got = sim.download_particles()
    sim.close()
    for f in ("position", "velocity", "density", "predicted_position"):
        assert np.array_equal(got[f], osim.p[f], equal_nan=True), f
    assert np.array_equal(got["grid"], osim.p["grid"])


def test_force_field_pushout_and_mouse(fluid, oracle):
    s, ts, p = scenes.uniform_box(64, tex=(32, 32))
    rng = np.random.default_rng(9)
    field = rng.integers(-4, 5, size=(32, 32, 2)).astype(np.float32)
    field[rng.random((32, 32)) < 0.4] = 0
    ts.mouse_state = 1
    ts.mouse_pos = (0.7, -1.1)
    ts.mouse_force_radius = 2.0
    p = p.copy()
    p["position"][17] = ts.mouse_pos  # dist == 0 -> NaN -> zeroed by the guard (compute.wgsl:113-116)
    p["predicted_position"][17] = ts.mouse_pos
    p["velocity"][17] = 0
    stats = _check(fluid, oracle, s, ts, p, field=field, n_ticks=2)
    assert stats["floats_bit_exact"], stats


def test_stale_start_indices_quirk(fluid, oracle):
    """the table is never cleared and the lowest occupied cell keeps its previous entry, from which the reference
    walks that cell (compute.wgsl:50, SURVEY.md §7 hard part 5). Exercise stale values inside, at the end of and
    beyond the first cell, and garbage everywhere else."""
    s, ts, p = scenes.uniform_box(48, tex=(16, 16))
    gw, gh = oracle.grid_dims(s)
    rng = np.random.default_rng(2)
    u = oracle.make_uniform(s, ts, 1)
    q = p.copy()
    oracle.predict_next_position(u, q)
    oracle.create_spatial_lookup(u, q)
    c0 = int(q["grid"].min())
    cnt0 = int((q["grid"] == c0).sum())
    assert cnt0 >= 2
    for stale in (1, cnt0 - 1, cnt0, cnt0 + 5, 10 ** 6, 0xFFFFFFFF):
        table = rng.integers(0, 4000, size=gw * gh, dtype=np.uint32)
        table[c0] = stale
        stats = _check(fluid, oracle, s, ts, p, table=table, n_ticks=2)
        assert stats["floats_bit_exact"], (stale, stats)


def test_table_carries_over_between_ticks(fluid, oracle):
    s, ts, p = scenes.dam_break(32, 96)
    osim, sim, _ = parity.run_pair(fluid, oracle, s, ts, p, n_ticks=25, debug_force=False)
    t = sim.download_start_indices()
    sim.close()
    assert np.array_equal(t, osim.table)
    occupied = np.unique(osim.p["grid"])
    stale = np.setdiff1d(np.nonzero(t)[0], occupied)
    assert stale.size > 0  # cells that emptied out keep their old entries, as in the reference


def test_upload_download_roundtrip_and_permutation(fluid, oracle):
    s, ts, p = scenes.uniform_box(80)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        assert np.array_equal(sim.download_particles(), p)
        sim.tick(ts)
        got = sim.download_particles()
        perm = sim.download_permutation()
        assert np.array_equal(np.sort(perm), np.arange(p.shape[0], dtype=np.uint32))
        keys = sim.download_unsorted_keys()
        assert np.array_equal(keys[perm], got["grid"])
        assert np.array_equal(perm, np.argsort(keys, kind="stable"))  # LSD radix == stable order


@pytest.mark.parametrize("name", ["uniform_48", "dambreak_24x96", "default_5000", "field_mouse_40"])
def test_golden_fixtures(fluid, name):
    """committed oracle vectors (tests/golden/make_golden.py); no oracle code runs in this test."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(GOLDEN, "make_golden.py"))
    mg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mg)
    with np.load(os.path.join(GOLDEN, name + ".npz")) as z:
        s, ts = mg.load_settings(z)
        with fluid.FluidSimulation.new(0, s) as sim:
            sim.upload_particles(np.ascontiguousarray(z["input"]))
            if "field" in z.files:
                sim.upload_force_field(z["field"])
            if "table_in" in z.files:
                sim.upload_start_indices(z["table_in"])
            for _ in range(int(z["ticks"][0])):
                sim.tick(ts)
            got = sim.download_particles()
            table = sim.download_start_indices()
        ref = z["output"]
        assert np.array_equal(got["grid"], ref["grid"])
        assert np.array_equal(table, z["table"])
        for f in ("position", "predicted_position", "velocity", "density"):
            assert np.allclose(got[f], ref[f], rtol=1e-5, atol=1e-6), f
            assert np.array_equal(got[f], ref[f]), f  # and in fact identical


def test_config1_default_scene_100_frames(fluid, oracle):
    """BASELINE.json config 1: the reference's default scene (src/main.rs:48-54: 100,000 particles on the init lattice of
    src/simulation.rs:147-163, 72 coincident pairs) stepped for 100 frames with the default tick settings
    (src/renderer.rs:374-388). (a) The whole trajectory: every step is bit-identical, so the states after 1, 50 and 100
    frames are too. (b) Per-step parity from the oracle's own states at ticks 1, 50 and 100 (identical input state ->
    one tick), all criteria of parity.compare_state."""
    s, ts = SimulationSettings(), TickSettings()
    osim = oracle.OracleSim(s, oracle.init_lattice(s))
    snaps = {}
    with fluid.FluidSimulation.new(0, s) as sim:  # create() builds the same lattice
        for t in range(1, 101):
            if t in (1, 50, 100):
                snaps[t] = (osim.p.copy(), osim.table.copy())
            osim.tick(ts)
            sim.tick(ts)
            if t in (1, 50, 100):
                got = sim.download_particles()
                assert np.array_equal(sim.download_start_indices(), osim.table), t
                for f in ("grid", "position", "predicted_position", "velocity", "density"):
                    assert np.array_equal(got[f], osim.p[f]), (t, f)
        assert sim.tick_count == 100
    for t, (state, table) in snaps.items():
        o2, s2, extras = parity.run_pair(fluid, oracle, s, ts, state, table=table, tick_count=t - 1)
        try:
            stats = parity.compare_state(o2, s2, extras, ts)
            assert stats["floats_bit_exact"], (t, stats)
        finally:
            s2.close()


def test_config2_hash_sort_table_1m_bit_exact(fluid, oracle):
    """BASELINE.json config 2: spatial hash + key sort + cell-start table, 1M particles, bit-exact."""
    for s, ts, p in (scenes.uniform_box(1000), scenes.dam_break(500, 2000)):
        osim = oracle.OracleSim(s, p, sort_mode=oracle.SORT_COUNTING)
        u = oracle.make_uniform(s, ts, 1)
        tmp = p.copy()
        oracle.predict_next_position(u, tmp)
        oracle.create_spatial_lookup(u, tmp)
        with fluid.FluidSimulation.new(0, s) as sim:
            sim.upload_particles(p)
            for k in range(2):  # second tick: table given the previous tick's table
                osim.tick(ts, stop_after=1)
                sim.tick_hash_sort_only(ts)
                if k == 0:
                    assert np.array_equal(sim.download_unsorted_keys(), tmp["grid"])
                got = sim.download_particles()
                assert np.array_equal(got["grid"], osim.p["grid"])
                assert np.array_equal(sim.download_start_indices(), osim.table)
                assert np.array_equal(got["position"], osim.p["position"])  # payload moved with its key, stable order


def test_config3_full_step_1m_dambreak(fluid, oracle):
    """BASELINE.json config 3: full SPH step on the 1M-particle dam-break, one tick from the initial state and one
    from the oracle's state after 3 ticks."""
    s, ts, p = scenes.dam_break(500, 2000)
    osim, sim, extras = parity.run_pair(fluid, oracle, s, ts, p, n_ticks=1, sort_mode=oracle.SORT_COUNTING, debug_force=False)
    stats = parity.compare_state(osim, sim, extras, ts)
    assert stats["floats_bit_exact"], stats
    for _ in range(2):
        osim.tick(ts)
    state, table = osim.p.copy(), osim.table.copy()
    sim.close()
    osim2, sim2, extras2 = parity.run_pair(fluid, oracle, s, ts, state, table=table, tick_count=3,
                                           sort_mode=oracle.SORT_COUNTING, debug_force=False)
    stats = parity.compare_state(osim2, sim2, extras2, ts)
    sim2.close()
    assert stats["floats_bit_exact"], stats


def test_16m_full_step_bit_exact(fluid, oracle):
    """16M particles, two ticks, every field of every particle against the oracle (the 64M scene of bench.py is checked
    the same way by scripts/verify_full_size.py; results: profiles/r1d_verify_64m.json on the counting path,
    profiles/r1c_verify_64m.json on the radix path)."""
    s, ts, p = scenes.uniform_box(4000)
    osim = oracle.OracleSim(s, p, sort_mode=oracle.SORT_COUNTING)  # stable counting sort == the stable radix order
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        del p
        for _ in range(2):
            sim.tick(ts)
            osim.tick(ts)
        got = sim.download_particles()
        table = sim.download_start_indices()
    for f in ("grid", "position", "predicted_position", "velocity", "density"):
        assert np.array_equal(np.ascontiguousarray(got[f]).view(np.uint32), np.ascontiguousarray(osim.p[f]).view(np.uint32)), f
    assert np.array_equal(table, osim.table)


@pytest.mark.parametrize("n_side", [4000])
def test_large_scene_properties(fluid, n_side):
    """size-independent properties of the sort / table / hash at 16M (no oracle involved)."""
    s, ts, p = scenes.uniform_box(n_side)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        del p
        sim.tick(ts)
        keys_u = sim.download_unsorted_keys()
        perm = sim.download_permutation()
        got = sim.download_particles()
        table = sim.download_start_indices()
    n = got.shape[0]
    assert (np.diff(got["grid"].astype(np.int64)) >= 0).all()  # sortedness
    assert np.array_equal(np.bincount(perm, minlength=n), np.ones(n, np.int64))  # a permutation
    assert np.array_equal(keys_u[perm], got["grid"])  # keys moved with their particles
    same = got["grid"][1:] == got["grid"][:-1]
    assert (np.diff(perm.astype(np.int64))[same] > 0).all()  # stable within a cell
    first = np.nonzero(~same)[0] + 1
    assert np.array_equal(table[got["grid"][first]], first.astype(np.uint32))  # cell-start table
    assert np.isfinite(got["density"]).all() and (got["density"] >= 0.1).all()
    gw = int(np.ceil(np.float32(s.size[0]) / np.float32(s.smoothing_radius))) + 2
    cx = np.floor((got["predicted_position"][:, 0] + np.float32(s.size[0]) * np.float32(0.5)) / np.float32(s.smoothing_radius)).astype(np.uint32) + 1
    cy = np.floor((got["predicted_position"][:, 1] + np.float32(s.size[1]) * np.float32(0.5)) / np.float32(s.smoothing_radius)).astype(np.uint32) + 1
    assert np.array_equal(cy * np.uint32(gw) + cx, got["grid"])  # key == hash(pred) (numpy fp32 division is IEEE)

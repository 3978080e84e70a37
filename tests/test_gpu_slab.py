"""Slab (multi-GPU) mode on ONE GPU: several ranks run one after another in one process (LocalDriver, device copies
instead of NCCL) and the gathered result must equal a single-GPU tick of the same state BIT FOR BIT: keys, sorted
order, the global start_indices table, densities, velocities, positions (SURVEY.md §8e "parity").
The same SlabRank phases run under torch.distributed/NCCL in bench.py --gpus N (tests/test_slab_protocol.py covers
the distributed driver with gloo on CPU)."""
import numpy as np
import pytest

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes

pytestmark = pytest.mark.gpu


def _run(fluid, s, ts, p, world, ticks, pad=4096, field=None, math_mode=0):
    from gpu_fluid_simulation_b200 import slab
    gw_gh = None
    ranks = [slab.SlabRank(0, s, r, world, capacity=p.shape[0], pad=pad, math_mode=math_mode) for r in range(world)]
    try:
        bands = slab.split_particles(p, s, ranks[0].bounds)
        order = np.concatenate(bands)  # the global array the slab run starts from
        for r, b in zip(ranks, bands):
            r.upload_band(b)
            if field is not None:
                r.sim.upload_force_field(field)
        drv = slab.LocalDriver(ranks)
        with fluid.FluidSimulation.new(0, s) as ref:
            ref.math_mode = math_mode
            ref.upload_particles(order)
            if field is not None:
                ref.upload_force_field(field)
            for _ in range(ticks):
                ref.tick(ts)
                drv.tick(ts)
            want = ref.download_particles()
            want_table = ref.download_start_indices()
            gw = ref.grid_w
        got = np.concatenate([r.download_band() for r in ranks])
        table = np.zeros_like(want_table)
        for r in ranks:
            lo, hi = r.owns_rows()
            t = r.sim.download_start_indices()
            table[lo * gw:hi * gw] = t[lo * gw:hi * gw]
        counts = [r.n for r in ranks]
        _run.last_paths = [r.sim.last_sort_path for r in ranks]
    finally:
        for r in ranks:
            r.close()
    return got, table, want, want_table, counts


def _assert_identical(got, table, want, want_table):
    assert got.shape == want.shape
    assert np.array_equal(got["grid"], want["grid"]), "gathered sorted keys differ from the single-GPU run"
    assert np.array_equal(table, want_table), "gathered start_indices differ"
    for f in ("predicted_position", "density", "velocity", "position"):
        assert np.array_equal(got[f], want[f]), f


@pytest.mark.parametrize("world", [2, 3, 4, 8])
def test_uniform_box_slabs(fluid, world):
    s, ts, p = scenes.uniform_box(160)
    got, table, want, want_table, counts = _run(fluid, s, ts, p, world, ticks=6)
    _assert_identical(got, table, want, want_table)
    assert min(counts) > 0
    # after the first tick (radix passes while the cell statistics are unknown) the bands count over their cells
    assert all(path == 2 for path in _run.last_paths), _run.last_paths


@pytest.mark.parametrize("world", [2, 5])
def test_dam_break_slabs_with_gravity_migration(fluid, world):
    s, ts, p = scenes.dam_break(48, 256)
    ts.gravity = (3.0, 9.8)
    got, table, want, want_table, counts = _run(fluid, s, ts, p, world, ticks=40)
    _assert_identical(got, table, want, want_table)


def test_default_scene_slabs_rng_uses_global_index(fluid, oracle):
    """coincident pairs: the xorshift seed is the GLOBAL sorted index (compute.wgsl:161)."""
    s = SimulationSettings()
    p = oracle.init_lattice(s)
    got, table, want, want_table, _ = _run(fluid, s, TickSettings(), p, 3, ticks=3)
    _assert_identical(got, table, want, want_table)


def test_fast_particles_cross_several_bands(fluid):
    """speeds up to the 500 clamp move particles ~21 cells per tick (compute.wgsl:118-122): multi-hop migration."""
    n = 6000
    s = SimulationSettings(particle_count=n, size=(12.0, 12.0), texture_size=(16, 16))
    rng = np.random.default_rng(5)
    p = np.zeros(n, dtype=fluid.PARTICLE_DTYPE)
    p["position"] = (rng.random((n, 2)) - 0.5) * 11.5
    p["velocity"] = rng.normal(size=(n, 2)) * 150
    p["predicted_position"] = p["position"]
    got, table, want, want_table, _ = _run(fluid, s, TickSettings(gravity=(0.0, 9.8)), p, 6, ticks=5)
    _assert_identical(got, table, want, want_table)


def test_empty_and_single_row_bands(fluid):
    """more ranks than occupied rows: some bands are empty, some one row tall."""
    n = 900
    s = SimulationSettings(particle_count=n, size=(8.0, 1.2), texture_size=(16, 16))  # 6 particle rows of cells
    rng = np.random.default_rng(8)
    p = np.zeros(n, dtype=fluid.PARTICLE_DTYPE)
    p["position"][:, 0] = (rng.random(n) - 0.5) * 7.8
    p["position"][:, 1] = (rng.random(n) - 0.5) * 0.5  # only the middle rows are occupied
    p["velocity"] = rng.normal(size=(n, 2))
    p["predicted_position"] = p["position"]
    got, table, want, want_table, counts = _run(fluid, s, TickSettings(), p, 8, ticks=4)
    _assert_identical(got, table, want, want_table)
    assert 0 in counts


def test_force_field_and_fast_math_slabs(fluid):
    s, ts, p = scenes.uniform_box(96, tex=(32, 32))
    rng = np.random.default_rng(9)
    field = rng.integers(-2, 3, size=(32, 32, 2)).astype(np.float32)
    got, table, want, want_table, _ = _run(fluid, s, ts, p, 4, ticks=3, field=field, math_mode=1)
    _assert_identical(got, table, want, want_table)

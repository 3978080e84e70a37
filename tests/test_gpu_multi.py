"""The multi-GPU step (fluid_multi_*, csrc/multi.inl) against a single FluidSim: after the same ticks from the same
uploaded array the bands, concatenated in rank order, must equal the single-GPU state BIT FOR BIT — sorted keys, the
order inside a cell, predicted positions, densities, velocities, positions, and the global start_indices table
(SURVEY.md §8e "parity"; the reference has one tick() per frame, src/simulation.rs:459).

Most cases run 2..8 ranks as bands that SHARE GPU 0 (FLUID_TRANSPORT_LOCAL: the same kernels and protocol, the exchange
windows pulled by a copy kernel instead of NCCL), so they need one GPU. The NCCL cases need >= 2 GPUs and skip
otherwise (the driver's GPU test box has one; `gpurun --gpus 2` runs them, and bench.py --gpus N checks the same
parity on its own before timing)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(fluid, s, ts, p, world, ticks, field=None, math_mode=0, table=None, devices=None, tick0=0, **opt):
    from gpu_fluid_simulation_b200 import multi
    devices = [0] * world if devices is None else devices
    if "transport" not in opt:
        opt["transport"] = multi.TRANSPORT_LOCAL
    with multi.MultiFluidSimulation.new(devices, s, **opt) as m, fluid.FluidSimulation.new(0, s) as ref:
        m.set_math_mode(math_mode)
        ref.math_mode = math_mode
        m.upload_particles(p)
        ref.upload_particles(p)
        if field is not None:
            m.upload_force_field(field)
            ref.upload_force_field(field)
        if table is not None:
            m.upload_start_indices(table)
            ref.upload_start_indices(table)
        if tick0:
            m.tick_count = tick0
            ref.tick_count = tick0
        for _ in range(ticks):
            ref.tick(ts)
            m.tick(ts)
        m.sync()
        want = ref.download_particles()
        want_table = ref.download_start_indices()
        got = m.download_particles()
        table_got = m.download_start_indices()
        counts = m.particle_counts()
        assert m.tick_count == ref.tick_count
    return got, table_got, want, want_table, counts


def _assert_identical(got, table, want, want_table):
    assert got.shape == want.shape
    assert np.array_equal(got["grid"], want["grid"]), "gathered sorted keys differ from the single-GPU run"
    for f in ("predicted_position", "density", "velocity", "position"):
        bad = np.nonzero((got[f] != want[f]).reshape(got.shape[0], -1).any(axis=1))[0]
        assert bad.size == 0, f"{f}: {bad.size} particles differ, first at sorted index {bad[:5]}"
    bad = np.nonzero(table != want_table)[0]
    assert bad.size == 0, f"start_indices: {bad.size} cells differ, first {bad[:5]}"


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_uniform_box_bands(fluid, world):
    s, ts, p = scenes.uniform_box(160)
    got, table, want, want_table, counts = _run(fluid, s, ts, p, world, ticks=6)
    _assert_identical(got, table, want, want_table)
    assert min(counts) > 0 and sum(counts) == s.particle_count


@pytest.mark.parametrize("world", [2, 5])
def test_dam_break_bands_with_gravity_migration(fluid, world):
    s, ts, p = scenes.dam_break(48, 256)
    ts.gravity = (3.0, 9.8)
    got, table, want, want_table, counts = _run(fluid, s, ts, p, world, ticks=40, window_near=16384)
    _assert_identical(got, table, want, want_table)


def test_default_scene_bands_rng_uses_global_index(fluid, oracle):
    """coincident pairs: the xorshift seed is the GLOBAL sorted index (compute.wgsl:161), and the order of two coincident
    particles inside their cell is the uploaded array's."""
    s = SimulationSettings()
    p = oracle.init_lattice(s)
    got, table, want, want_table, _ = _run(fluid, s, TickSettings(), p, 3, ticks=3)
    _assert_identical(got, table, want, want_table)


def test_shuffled_upload_order_is_the_tie_order(fluid):
    """the uploaded array is in arbitrary order (first tick after an upload carries explicit origins)."""
    s, ts, p = scenes.uniform_box(96)
    rng = np.random.default_rng(3)
    p = p[rng.permutation(p.shape[0])]
    p[100]["position"] = p[7]["position"]  # two coincident particles far apart in the array
    p[100]["predicted_position"] = p[7]["predicted_position"]
    got, table, want, want_table, _ = _run(fluid, s, ts, p, 4, ticks=3, window_near=16384)
    _assert_identical(got, table, want, want_table)


def test_fast_particles_cross_several_bands(fluid):
    """speeds up to the 500 clamp move particles ~21 cells per tick (compute.wgsl:118-122): jumps over whole bands."""
    n = 6000
    s = SimulationSettings(particle_count=n, size=(12.0, 12.0), texture_size=(16, 16))
    rng = np.random.default_rng(5)
    p = np.zeros(n, dtype=fluid.PARTICLE_DTYPE)
    p["position"] = (rng.random((n, 2)) - 0.5) * 11.5
    p["velocity"] = rng.normal(size=(n, 2)) * 150
    p["predicted_position"] = p["position"]
    got, table, want, want_table, _ = _run(fluid, s, TickSettings(gravity=(0.0, 9.8)), p, 6, ticks=5, window_far=8192, capacity=n)
    _assert_identical(got, table, want, want_table)


def test_empty_and_single_row_bands(fluid):
    """more ranks than occupied rows: some bands are empty, some one row tall (ghost rows then come from two ranks away)."""
    n = 900
    s = SimulationSettings(particle_count=n, size=(8.0, 1.2), texture_size=(16, 16))  # 6 particle rows of cells
    rng = np.random.default_rng(8)
    p = np.zeros(n, dtype=fluid.PARTICLE_DTYPE)
    p["position"][:, 0] = (rng.random(n) - 0.5) * 7.8
    p["position"][:, 1] = (rng.random(n) - 0.5) * 0.5  # only the middle rows are occupied
    p["velocity"] = rng.normal(size=(n, 2))
    p["predicted_position"] = p["position"]
    got, table, want, want_table, counts = _run(fluid, s, TickSettings(), p, 8, ticks=4, window_far=4096, capacity=n)
    _assert_identical(got, table, want, want_table)
    assert 0 in counts


def test_force_field_and_fast_math_bands(fluid):
    s, ts, p = scenes.uniform_box(96, tex=(32, 32))
    rng = np.random.default_rng(9)
    field = rng.integers(-2, 3, size=(32, 32, 2)).astype(np.float32)
    got, table, want, want_table, _ = _run(fluid, s, ts, p, 4, ticks=3, field=field, math_mode=1, window_far=8192, capacity=96 * 96)
    _assert_identical(got, table, want, want_table)


def test_stale_table_quirk_bands(fluid):
    """the lowest occupied cell is walked from its stale start_indices entry (compute.wgsl:50; never-cleared table,
    src/simulation.rs:204-209): garbage in the uploaded table must act exactly as on one GPU, also when the band that
    holds that cell is not the one that computes its neighbours."""
    s, ts, p = scenes.uniform_box(64)
    gw = int(np.ceil(np.float32(s.size[0]) / np.float32(s.smoothing_radius))) + 2
    rng = np.random.default_rng(11)
    table = rng.integers(0, p.shape[0] + 50, size=gw * gw).astype(np.uint32)
    for world in (2, 3):
        got, tab, want, want_table, _ = _run(fluid, s, ts, p, world, ticks=3, table=table, tick0=17)
        _assert_identical(got, tab, want, want_table)


def test_changing_delta_drops_the_speculated_key_pass(fluid):
    """the reference's delta is a setting the user can drag while the simulation runs (src/renderer.rs:642-644), so it can differ
    between two ticks: the key pass that the force pass ran ahead with the old delta — keys, histogram, copies for the other ranks — must be dropped, not added to."""
    from gpu_fluid_simulation_b200 import multi
    s, ts, p = scenes.dam_break(48, 256)
    ts.gravity = (3.0, 9.8)
    deltas = [1 / 120, 1 / 120, 1 / 90, 1 / 144, 1 / 144, 1 / 60, 1 / 120]
    for world in (2, 5):
        with multi.MultiFluidSimulation.new([0] * world, s, transport=multi.TRANSPORT_LOCAL, window_near=16384) as m, \
                fluid.FluidSimulation.new(0, s) as ref:
            m.upload_particles(p)
            ref.upload_particles(p)
            for d in deltas:
                ts.delta = d
                ref.tick(ts)
                m.tick(ts)
            m.sync()
            _assert_identical(m.download_particles(), m.download_start_indices(), ref.download_particles(), ref.download_start_indices())


def test_custom_and_balanced_row_bounds(fluid):
    from gpu_fluid_simulation_b200 import multi
    s, ts, p = scenes.dam_break(40, 200)
    b = multi.balanced_bounds(s, p, 4)
    assert b[0] == 0 and b[-1] == int(np.ceil(np.float32(s.size[1]) / np.float32(s.smoothing_radius))) + 2
    assert all(b[i] <= b[i + 1] for i in range(4))
    got, table, want, want_table, counts = _run(fluid, s, ts, p, 4, ticks=5, row_bounds=b, window_near=8192)
    _assert_identical(got, table, want, want_table)
    assert max(counts) - min(counts) <= 2 * 40 * 2 + 80  # within a couple of lattice rows of each other
    # deliberately lopsided bounds
    gh = b[-1]
    got, table, want, want_table, counts = _run(fluid, s, ts, p, 3, ticks=4, row_bounds=[0, 3, gh - 2, gh], window_near=8192, capacity=p.shape[0])
    _assert_identical(got, table, want, want_table)


def test_window_overflow_is_reported(fluid):
    from gpu_fluid_simulation_b200 import multi
    s, ts, p = scenes.uniform_box(128)
    with multi.MultiFluidSimulation.new([0, 0], s, transport=multi.TRANSPORT_LOCAL, window_near=64) as m:
        m.upload_particles(p)
        m.tick(ts)
        with pytest.raises(fluid.FluidError, match="overflow"):
            m.sync()


def test_many_ticks_without_host_reads(fluid):
    """tick_n enqueues whole ticks back to back (no size ever reaches the host); state still identical after 60 ticks."""
    s, ts, p = scenes.uniform_box(128)
    from gpu_fluid_simulation_b200 import multi
    with multi.MultiFluidSimulation.new([0, 0, 0], s, transport=multi.TRANSPORT_LOCAL) as m, fluid.FluidSimulation.new(0, s) as ref:
        m.upload_particles(p)
        ref.upload_particles(p)
        m.tick_n(ts, 60)
        ref.tick_n(ts, 60)
        _assert_identical(m.download_particles(), m.download_start_indices(), ref.download_particles(), ref.download_start_indices())


def test_pipelined_band_transfers(fluid):
    """fluid_multi_upload_band_async / _download_band_async on a one-rank handle: same bytes as the synchronous calls and as
    a plain FluidSim, with nothing read back by the host between upload and download."""
    import torch
    from gpu_fluid_simulation_b200 import multi
    s, ts, p = scenes.uniform_box(200)
    n = p.shape[0]
    src = torch.empty((n, 8), dtype=torch.float32).pin_memory()
    srcv = src.numpy().view(fluid.PARTICLE_DTYPE).reshape(-1)
    srcv[:] = p
    dst = torch.empty((n + 64, 8), dtype=torch.float32).pin_memory()
    dstv = dst.numpy().view(fluid.PARTICLE_DTYPE).reshape(-1)
    info = torch.zeros(2, dtype=torch.int32).pin_memory()
    with multi.MultiFluidSimulation.new_rank(0, 1, None, 0, s, capacity=n + 64) as m, fluid.FluidSimulation.new(0, s) as ref:
        for _ in range(3):  # every round replaces the state, as bench.py's e2e loop does
            m.upload_band_async(srcv)
            m.tick_n(ts, 2)
            m.download_band_async(dstv, info.numpy().view(np.uint32))
        m.sync_transfers()
        ref.upload_particles(p)
        ref.tick_n(ts, 2)
        want = ref.download_particles()
        assert int(info[0]) == n and int(info[1]) == 0
        assert np.array_equal(dstv[:n], want)
        got, goff = m.download_band()
        assert goff == 0 and np.array_equal(got, want)


def _gpus():
    import torch
    return torch.cuda.device_count()


def test_nccl_single_process_two_gpus(fluid):
    """one process drives two GPUs through NCCL (ncclCommInitAll) and through peer-memory pulls."""
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from gpu_fluid_simulation_b200 import multi
    s, ts, p = scenes.uniform_box(512)
    for transport in (multi.TRANSPORT_NCCL, multi.TRANSPORT_LOCAL):
        got, table, want, want_table, _ = _run(fluid, s, ts, p, 2, ticks=5, devices=[0, 1], transport=transport)
        _assert_identical(got, table, want, want_table)
    s, ts, p = scenes.dam_break(128, 512)
    got, table, want, want_table, _ = _run(fluid, s, ts, p, 2, ticks=30, devices=[0, 1], transport=multi.TRANSPORT_NCCL, window_near=32768)
    _assert_identical(got, table, want, want_table)


@pytest.mark.parametrize("transport", ["peer", "nccl"])
@pytest.mark.parametrize("dam", [False, True])
def test_one_rank_per_process(fluid, transport, dam):
    """torchrun, one rank per GPU: scripts/multi_parity.py gathers the bands on rank 0 and compares with one GPU. The
    exchange is the CUDA-IPC pull over NVLink by default (FLUID_TRANSPORT_PEER) and NCCL send/recv with FLUID_MULTI_NO_PEER=1."""
    n = _gpus()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29631", os.path.join(ROOT, "scripts", "multi_parity.py"), "--n-side", "1024", "--ticks", "7"] + (["--dam"] if dam else [])
    env = dict(os.environ, FLUID_MULTI_NO_PEER="1" if transport == "nccl" else "0")
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert "fields_differing 0" in res.stdout
    assert f"transport {transport}" in res.stdout

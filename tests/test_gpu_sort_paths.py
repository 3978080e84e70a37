"""The two key-order paths of the single-GPU tick — counting sort over the cells (kernels_count.cuh, default) and the
onesweep radix passes (radix_sort.cuh) — must produce the same stable order, hence identical bits everywhere, and both
must match the oracle (sort.wgsl:27-51 + compute.wgsl:45-56 restated: sorted keys and start_indices bit-exact)."""
import numpy as np
import pytest

import parity
from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes

pytestmark = pytest.mark.gpu
AUTO, RADIX, COUNTING = 0, 1, 2


def _crowded(oracle, n=700):
    s = SimulationSettings(particle_count=n, size=(4.0, 4.0), texture_size=(8, 8))
    rng = np.random.default_rng(1)
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = 0.05 + rng.random((n, 2)) * 0.1
    p["position"][100:140] = p["position"][0]
    p["predicted_position"] = p["position"]
    return s, TickSettings(), p


def _scene(oracle, name):
    if name == "uniform":
        return scenes.uniform_box(200)
    if name == "dam":
        return scenes.dam_break(64, 256)
    if name == "default":
        s = SimulationSettings()
        return s, TickSettings(), oracle.init_lattice(s)
    if name == "shuffled":  # arbitrary input order: no runs along the lanes, scattered slots
        s, ts, p = scenes.uniform_box(150)
        return s, ts, p[np.random.default_rng(5).permutation(p.shape[0])]
    if name == "crowded":
        return _crowded(oracle)
    raise KeyError(name)


@pytest.mark.parametrize("name", ["uniform", "dam", "default", "shuffled", "crowded"])
@pytest.mark.parametrize("mode", [RADIX, COUNTING])
def test_forced_path_matches_oracle(fluid, oracle, name, mode):
    s, ts, p = _scene(oracle, name)
    osim, sim, extras = parity.run_pair(fluid, oracle, s, ts, p, n_ticks=3, gpu_sort_mode=mode)
    try:
        assert sim.sort_mode == mode and sim.last_sort_path == mode
        stats = parity.compare_state(osim, sim, extras, ts)
        assert stats["floats_bit_exact"], stats
    finally:
        sim.close()


@pytest.mark.parametrize("name", ["uniform", "dam", "default", "shuffled", "crowded"])
def test_paths_give_identical_bits(fluid, oracle, name):
    s, ts, p = _scene(oracle, name)
    out = {}
    for mode in (RADIX, COUNTING, AUTO):
        with fluid.FluidSimulation.new(0, s) as sim:
            sim.sort_mode = mode
            sim.upload_particles(p)
            sim.tick(ts)
            first_path = sim.last_sort_path
            for _ in range(3):
                sim.tick(ts)
            out[mode] = (sim.download_particles(), sim.download_permutation(), sim.download_start_indices(),
                         sim.download_unsorted_keys(), first_path)
    for mode in (COUNTING, AUTO):
        for a, b in zip(out[RADIX][:4], out[mode][:4]):
            assert np.array_equal(a, b), (name, mode)
    # AUTO, first tick after the upload: counting unless a cell is crowded (> 64 particles)
    assert out[AUTO][4] == (RADIX if name == "crowded" else COUNTING)
    perm, keys = out[COUNTING][1], out[COUNTING][3]
    assert np.array_equal(perm, np.argsort(keys, kind="stable"))


@pytest.mark.parametrize("n", [2, 3, 31, 32, 33, 255, 257, 4097, 8191])
def test_counting_path_ragged_counts(fluid, oracle, n):
    rng = np.random.default_rng(n)
    s = SimulationSettings(particle_count=n, size=(6.0, 4.0), texture_size=(16, 16))
    p = np.zeros(n, dtype=oracle.PARTICLE_DTYPE)
    p["position"] = (rng.random((n, 2)) - 0.5) * np.array([6.0, 4.0]) * 0.98
    p["velocity"] = rng.normal(size=(n, 2)) * 3
    p["predicted_position"] = p["position"]
    osim, sim, extras = parity.run_pair(fluid, oracle, s, TickSettings(gravity=(0.0, 9.8)), p, n_ticks=2, gpu_sort_mode=COUNTING)
    try:
        stats = parity.compare_state(osim, sim, extras, TickSettings(gravity=(0.0, 9.8)))
        assert stats["floats_bit_exact"], stats
    finally:
        sim.close()


def test_auto_switches_when_cells_get_crowded_and_back(fluid, oracle):
    """AUTO follows the largest cell population: a crowded upload takes the radix passes, a spread-out one counting;
    hash-sort-only ticks (BASELINE config 2) go through the same switch."""
    s, ts, p = _crowded(oracle)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        sim.tick(ts)
        assert sim.last_sort_path == RADIX
        q = p.copy()
        rng = np.random.default_rng(3)
        q["position"] = (rng.random((p.shape[0], 2)) - 0.5) * 3.9
        q["predicted_position"] = q["position"]
        sim.upload_particles(q)
        sim.tick_hash_sort_only(ts)
        assert sim.last_sort_path == COUNTING
        got = sim.download_particles()
        keys = sim.download_unsorted_keys()
        assert np.array_equal(got["grid"], np.sort(keys))
        assert np.array_equal(sim.download_permutation(), np.argsort(keys, kind="stable"))
        assert sim.tick_count == 2 and sim.uniform().frame_time == 2


def test_speculated_keys_survive_changes_between_ticks(fluid, oracle):
    """the force pass emits the next tick's keys and cell counts assuming an unchanged delta and an untouched state; a
    tick with another delta, an upload, a hash-sort-only tick or a mode switch in between must drop them. Every step is
    checked against the oracle (bit-exact) and against the radix path, which never speculates."""
    s, ts, p = scenes.dam_break(48, 160)
    import dataclasses
    deltas = [ts.delta, ts.delta, ts.delta * 0.5, ts.delta * 0.5, ts.delta, ts.delta * 2.0, ts.delta, ts.delta]
    osim = oracle.OracleSim(s, p)
    sims = {m: fluid.FluidSimulation.new(0, s) for m in (AUTO, RADIX)}
    try:
        for m, sim in sims.items():
            sim.sort_mode = m
            sim.upload_particles(p)
        for k, d in enumerate(deltas):
            tk = dataclasses.replace(ts, delta=d) if dataclasses.is_dataclass(ts) else type(ts)(**{**ts.__dict__, "delta": d})
            if k == 5:  # replace the state between two ticks: the same particles, shuffled
                state = sims[RADIX].download_particles()
                perm = np.random.default_rng(9).permutation(state.shape[0])
                table = sims[RADIX].download_start_indices()
                osim.p = state[perm].copy()
                for sim in sims.values():
                    sim.upload_particles(state[perm])
                    sim.upload_start_indices(table)
            osim.tick(tk)
            outs = {}
            for m, sim in sims.items():
                sim.tick(tk)
                outs[m] = (sim.download_particles(), sim.download_start_indices(), sim.download_unsorted_keys())
            for a, b in zip(outs[AUTO], outs[RADIX]):
                assert np.array_equal(a, b), k
            got = outs[AUTO][0]
            assert np.array_equal(outs[AUTO][1], osim.table), k
            for f in ("grid", "position", "predicted_position", "velocity", "density"):
                assert np.array_equal(got[f], osim.p[f]), (k, f)
            if k == 3:  # a hash-sort-only tick in between (reorders the state, runs no force pass)
                for sim in sims.values():
                    sim.tick_hash_sort_only(tk)
                osim.tick(tk, stop_after=1)
    finally:
        for sim in sims.values():
            sim.close()

"""N > 1 path on CPU: the multi-GPU step's protocol (one exchange of ghost rows + migrants per tick, fixed-size windows,
origins, migration matrix -> global offsets; csrc/kernels_band.cuh, multi.inl) restated in tests/cpu_band_model.py and
run under torch.distributed / gloo with world_size 2 and 3, the oracle doing each rank's arithmetic. The gathered bands
must equal a single-domain oracle run bit for bit — which is what the CUDA path is held to on the GPU
(tests/test_gpu_multi.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def test_row_bounds_and_band_generation():
    from gpu_fluid_simulation_b200 import multi_bench
    import cpu_band_model as M
    for gh in (3, 5, 28, 267, 502, 4002):
        for world in (1, 2, 3, 8):
            b = multi_bench.default_row_bounds(gh, world)
            assert b == M.default_bounds(gh, world)
            assert len(b) == world + 1 and b[0] == 0 and b[-1] == gh and all(b[i] <= b[i + 1] for i in range(world))
    assert multi_bench.default_row_bounds(4002, 8) == [0, 501, 1001, 1501, 2001, 2501, 3001, 3501, 4002]
    # the bands the ranks generate locally are contiguous ranges of the scene's array: concatenated, they ARE the scene
    for kind, arg, full in (("uniform", 120, scenes.uniform_box(120)[2]), ("dam", (30, 120), scenes.dam_break(30, 120)[2])):
        s = multi_bench.scene_dims(kind, arg)[0]
        b = multi_bench.default_row_bounds(multi_bench.grid_h_of(s), 4)
        bands = [multi_bench.make_band(kind, arg, b, r) for r in range(4)]
        assert np.array_equal(np.concatenate(bands), full)
        h = np.float32(s.smoothing_radius)
        for r, band in enumerate(bands):
            row = np.floor((band["position"][:, 1] + np.float32(s.size[1]) * np.float32(0.5)) / h).astype(int) + 1
            assert row.min() >= b[r] and row.max() < b[r + 1]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _scene(name, O):
    if name == "uniform":
        s, ts, p = scenes.uniform_box(72)
    elif name == "dam":
        s, ts, p = scenes.dam_break(24, 120)
        ts.gravity = (4.0, 9.8)
    elif name == "fast":
        n = 1500
        s = SimulationSettings(particle_count=n, size=(9.0, 9.0), texture_size=(16, 16))
        rng = np.random.default_rng(5)
        p = np.zeros(n, dtype=O.PARTICLE_DTYPE)
        p["position"] = (rng.random((n, 2)) - 0.5) * 8.6
        p["velocity"] = rng.normal(size=(n, 2)) * 120  # many cross several bands per tick
        p["predicted_position"] = p["position"]
        ts = TickSettings(gravity=(0.0, 9.8))
    elif name == "shuffled":
        s, ts, p = scenes.uniform_box(60)
        p = p[np.random.default_rng(3).permutation(p.shape[0])]
        p[100]["position"] = p[7]["position"]
    else:  # the reference's sheared lattice: coincident pairs -> RNG seeded with the GLOBAL sorted index
        s = SimulationSettings(particle_count=6000, size=(13.0, 13.0), texture_size=(16, 16))
        ts, p = TickSettings(), O.init_lattice(s)
    return s, ts, p


def _worker(rank, world, port, scene, ticks, stale_table):
    import torch.distributed as dist
    import cpu_band_model as M
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s, ts, p = _scene(scene, O)
        r = M.BandRankModel(s, rank, world, window_near=p.shape[0], window_far=p.shape[0])
        # any split works (the first tick moves every particle to the rank that owns its row): deal the array out in chunks
        cuts = [p.shape[0] * k // world for k in range(world + 1)]
        table0 = None
        if stale_table:
            table0 = np.random.default_rng(11).integers(0, p.shape[0] + 50, size=r.gw * r.gh).astype(np.uint32)
            r.table[:] = table0
        r.upload(p[cuts[rank]:cuts[rank + 1]], np.arange(cuts[rank], cuts[rank + 1]), [cuts[k + 1] - cuts[k] for k in range(world)])
        for _ in range(ticks):
            out = r.key_pass(ts)
            inbox = M.exchange(dist, rank, world, out, r.cap)
            r.recv_sort(inbox)
            r.density_force()
        gathered = [None] * world if rank == 0 else None
        dist.gather_object((r.sorted, r.table, r.bounds[rank], r.bounds[rank + 1], r.error, r.goff), gathered, dst=0)
        if rank == 0:
            ref = O.OracleSim(s, p)
            if table0 is not None:
                ref.table[:] = table0
            for _ in range(ticks):
                ref.tick(ts)
            assert all(g[4] == 0 for g in gathered), "a rank latched an error"
            got = np.concatenate([g[0] for g in gathered])
            tab = np.zeros_like(ref.table)
            for g in gathered:
                tab[g[2] * r.gw:g[3] * r.gw] = g[1][g[2] * r.gw:g[3] * r.gw]
            assert got.shape == ref.p.shape
            assert [g[5] for g in gathered] == list(np.cumsum([0] + [g[0].shape[0] for g in gathered])[:-1]), "global offsets"
            assert np.array_equal(got["grid"], ref.p["grid"])
            assert np.array_equal(tab, ref.table)
            for f in ("predicted_position", "density", "velocity", "position"):
                assert np.array_equal(got[f], ref.p[f]), f
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,scene,ticks,stale", [(2, "uniform", 4, False), (3, "dam", 10, False), (2, "default", 3, False),
                                                     (3, "fast", 3, False), (2, "shuffled", 3, False), (3, "uniform", 3, True)])
def test_band_protocol_over_gloo_matches_single_domain(oracle, world, scene, ticks, stale):
    mp.spawn(_worker, args=(world, _free_port(), scene, ticks, stale), nprocs=world, join=True)


def _overflow_worker(rank, world, port):
    import torch.distributed as dist
    import cpu_band_model as M
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s, ts, p = _scene("uniform", O)
        r = M.BandRankModel(s, rank, world, window_near=16, window_far=16)
        cuts = [p.shape[0] * k // world for k in range(world + 1)]
        r.upload(p[cuts[rank]:cuts[rank + 1]], np.arange(cuts[rank], cuts[rank + 1]), [cuts[k + 1] - cuts[k] for k in range(world)])
        out = r.key_pass(ts)
        r.recv_sort(M.exchange(dist, rank, world, out, r.cap))
        assert r.error & M.ERR_WINDOW, "an overflowing window must latch the error bit on sender and receiver"
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_window_overflow_latches_an_error(oracle):
    mp.spawn(_overflow_worker, args=(2, _free_port()), nprocs=2, join=True)

"""N > 1 path on CPU: the product's DistDriver (migration all_to_all, halo P2P, metadata all_gather, ordering rules)
runs under torch.distributed/gloo with world_size 2 and 3, with the oracle doing the arithmetic of each rank
(tests/cpu_slab_backend.py). The gathered bands must equal a single-domain oracle run bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes
from gpu_fluid_simulation_b200 import slab


def test_row_bounds_cover_the_grid():
    for gh in (3, 5, 28, 267, 502, 4002):
        for world in (1, 2, 3, 8):
            b = slab.row_bounds(gh, world)
            assert len(b) == world + 1 and b[0] == 0 and b[-1] == gh
            assert all(b[i] <= b[i + 1] for i in range(world))
    assert slab.row_bounds(4002, 8) == [0, 501, 1001, 1501, 2001, 2501, 3001, 3501, 4002]


def test_band_generation_partitions_the_scene():
    n_side, world = 120, 4
    s, ts, full = scenes.uniform_box(n_side)
    bands = [slab.band_of_uniform_box(n_side, r, world)[2] for r in range(world)]
    assert np.array_equal(np.concatenate(bands), full)  # contiguous ranges of the full array, no rank holds it all
    h = np.float32(s.smoothing_radius)
    gh = int(np.ceil(np.float32(s.size[1]) / h)) + 2
    b = slab.row_bounds(gh, world)
    for r, band in enumerate(bands):
        row = np.floor((band["position"][:, 1] + np.float32(s.size[1]) * np.float32(0.5)) / h).astype(int) + 1
        assert row.min() >= b[r] and row.max() < b[r + 1]
    assert [x.shape[0] for x in slab.split_particles(full, s, b)] == [x.shape[0] for x in bands]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, scene, ticks, resolved):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch.distributed as dist
    from cpu_slab_backend import CpuSlabRank
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        if scene == "uniform":
            s, ts, p = scenes.uniform_box(72)
        elif scene == "dam":
            s, ts, p = scenes.dam_break(24, 120)
            ts.gravity = (4.0, 9.8)
        elif scene == "fast":
            n = 1500
            s = SimulationSettings(particle_count=n, size=(9.0, 9.0), texture_size=(16, 16))
            rng = np.random.default_rng(5)
            p = np.zeros(n, dtype=O.PARTICLE_DTYPE)
            p["position"] = (rng.random((n, 2)) - 0.5) * 8.6
            p["velocity"] = rng.normal(size=(n, 2)) * 120  # many cross several bands per tick
            p["predicted_position"] = p["position"]
            ts = TickSettings(gravity=(0.0, 9.8))
        else:  # the reference's sheared lattice: coincident pairs -> RNG seeded with the GLOBAL sorted index
            s = SimulationSettings(particle_count=6000, size=(13.0, 13.0), texture_size=(16, 16))
            ts, p = TickSettings(), O.init_lattice(s)
        r = CpuSlabRank(s, rank, world)
        bands = slab.split_particles(p, s, r.bounds)
        r.upload_band(bands[rank])
        drv = slab.DistDriver(r, device_resolved=resolved)
        for _ in range(ticks):
            drv.tick(ts)
        lo, hi = r.owns_rows()
        gathered = [None] * world if rank == 0 else None
        dist.gather_object((r.download_band(), r.table, lo, hi), gathered, dst=0)
        if rank == 0:
            ref = O.OracleSim(s, np.concatenate(bands))
            for _ in range(ticks):
                ref.tick(ts)
            got = np.concatenate([g[0] for g in gathered])
            tab = np.zeros_like(ref.table)
            for g in gathered:
                tab[g[2] * r.gw:g[3] * r.gw] = g[1][g[2] * r.gw:g[3] * r.gw]
            assert got.shape == ref.p.shape
            assert np.array_equal(got["grid"], ref.p["grid"])
            assert np.array_equal(tab, ref.table)
            for f in ("predicted_position", "density", "velocity", "position"):
                assert np.array_equal(got[f], ref.p[f]), f
            assert sum(int(g[0].shape[0] > 0) for g in gathered) >= 2
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("resolved", [False, True], ids=["host-resolved", "device-resolved"])
@pytest.mark.parametrize("world,scene,ticks", [(2, "uniform", 4), (3, "dam", 10), (2, "default", 3), (3, "fast", 3)])
def test_dist_driver_over_gloo_matches_single_domain(oracle, world, scene, ticks, resolved):
    """both exchange protocols of DistDriver: metadata read back by the host between sort and cells (two host reads per
    tick), and the device-resolved one (fixed-size halo windows, one host read per tick)."""
    mp.spawn(_worker, args=(world, _free_port(), scene, ticks, resolved), nprocs=world, join=True)

"""Pipelined host<->device transfers (fluid_sim_upload_particles_async / _download_particles_async): same bytes as the
synchronous calls, with inputs and outputs streaming through pinned host buffers."""
import numpy as np
import pytest

from gpu_fluid_simulation_b200 import scenes

pytestmark = pytest.mark.gpu


def _inputs(n_side, count):
    s, ts, p0 = scenes.uniform_box(n_side)
    rng = np.random.default_rng(11)
    states = []
    for k in range(count):
        p = p0[rng.permutation(p0.shape[0])].copy() if k else p0.copy()
        p["velocity"] = rng.normal(size=p["velocity"].shape).astype(np.float32) * (k + 1)
        states.append(p)
    return s, ts, states


@pytest.mark.parametrize("reuse_out", [False, True])
def test_pipelined_transfers_match_synchronous(fluid, reuse_out):
    import torch
    s, ts, states = _inputs(150, 5)
    n = states[0].shape[0]
    want = []
    with fluid.FluidSimulation.new(0, s) as sim:
        for st in states:
            sim.upload_particles(st)
            sim.tick(ts)
            want.append(sim.download_particles())
    ins = [torch.from_numpy(st.view(np.uint8).reshape(-1).copy()).pin_memory() for st in states]
    outs = [torch.zeros(n * 32, dtype=torch.uint8).pin_memory() for _ in range(2 if reuse_out else len(states))]
    got = [None] * len(states)
    with fluid.FluidSimulation.new(0, s) as sim:
        for k in range(len(states)):
            slot = k % len(outs)
            if k >= len(outs):
                sim.sync_transfers()  # the consumer reads a buffer (step k - len(outs)) before handing it back
                got[k - len(outs)] = outs[slot].numpy().copy()
            sim.upload_particles_async(ins[k])
            sim.tick(ts)
            sim.download_particles_async(outs[slot])
        sim.sync_transfers()
        for k in range(max(0, len(states) - len(outs)), len(states)):
            got[k] = outs[k % len(outs)].numpy().copy()
    assert len(got) == len(want)
    for k, (g, w) in enumerate(zip(got, want)):
        assert np.array_equal(g, w.view(np.uint8).reshape(-1)), f"step {k}"


def test_async_and_sync_calls_interleave(fluid):
    import torch
    s, ts, states = _inputs(64, 2)
    n = states[0].shape[0]
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(states[0]); sim.tick(ts); a = sim.download_particles()
        buf = torch.from_numpy(states[0].view(np.uint8).reshape(-1).copy()).pin_memory()
        out = torch.zeros(n * 32, dtype=torch.uint8).pin_memory()
        sim.tick_count = 0
        sim.upload_particles_async(buf); sim.tick(ts)
        b = sim.download_particles()          # synchronous download after an async upload
        sim.download_particles_async(out); sim.sync_transfers()
    assert np.array_equal(a, b)
    assert np.array_equal(out.numpy(), a.view(np.uint8).reshape(-1))

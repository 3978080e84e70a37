"""headless fluid render (fluid_shader.wgsl). CPU: sanity of the oracle's restatement.
GPU: the CUDA kernel against it. exp/log are library functions on both sides (libm vs CUDA, <= 2 ulp), so the
bar is a stated tolerance: 2e-5 absolute on every colour channel, and the blended UNORM8 image within 1 level."""
import numpy as np
import pytest

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, field as F, scenes

TOL = 2e-5


def test_oracle_render_sanity(oracle):
    s, ts, p = scenes.dam_break(24, 96, tex=(64, 64))
    sim = oracle.OracleSim(s, p)
    for _ in range(2):
        sim.tick(ts)
    inv = F.inverse_projection(s.size)
    img = oracle.render_fluid(sim.last_uniform, sim.p, sim.table, inv, 96, 96)
    assert img.shape == (96, 96, 4) and np.isfinite(img).all()
    a = img[..., 3]
    assert a.min() >= 0.0 and a.max() <= 1.0
    # the block of fluid is on the left (x < 0 half): opaque there, empty on the far right
    assert a[:, :20].mean() > 0.9 and a[:, 70:].max() == 0.0
    # where alpha is 0 the interior factor is 0, so the base colour vanishes (only the edge term could remain, and
    # smoothstep(0.7,1,d) is 0 whenever smoothstep(0.5,1.5,d) is 0 ... for d <= 0.5)
    assert (img[a == 0.0][:, :3] == 0.0).all()
    # an empty particle set renders to transparent black
    empty = oracle.render_fluid(sim.last_uniform, sim.p, np.full_like(sim.table, 0xFFFFFFFF), inv, 16, 16)
    assert (empty == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("scene,dims", [("dam", (960, 540)), ("uniform", (333, 201)), ("default", (256, 256))])
def test_gpu_render_matches_oracle(fluid, oracle, scene, dims):
    if scene == "dam":
        s, ts, p = scenes.dam_break(64, 256, tex=(128, 128))
    elif scene == "uniform":
        s, ts, p = scenes.uniform_box(96, tex=(128, 128))
    else:
        s, ts, p = SimulationSettings(particle_count=5000, size=(16.0, 16.0), texture_size=(64, 64)), TickSettings(), None
    w, h = dims
    inv = F.inverse_projection((s.size[0] * 1.1, s.size[1] * 1.1))  # view slightly larger than the box: off-grid pixels too
    with fluid.FluidSimulation.new(0, s) as sim:
        if p is not None:
            sim.upload_particles(p)
        for _ in range(3):
            sim.tick(ts)
        got, got8, ms = F.render(sim, inv, w, h, want_f32=True, want_rgba8=True, timed=True)
        particles = sim.download_particles()
        table = sim.download_start_indices()
        u = sim.uniform()
    ref = oracle.render_fluid(u, particles, table, inv, w, h)
    assert ms > 0
    assert np.isfinite(got).all()
    err = np.abs(got - ref)
    assert err.max() <= TOL, f"max abs colour error {err.max()} at {np.unravel_index(err.argmax(), err.shape)}"
    assert (ref[..., 3] > 0).mean() > 0.05  # the scene is actually in view
    blend = np.clip(ref[..., :3] * ref[..., 3:4], 0, 1) * 255.0
    assert np.abs(got8[..., :3].astype(np.float32) - blend).max() <= 1.0
    assert (got8[..., 3] == 255).all()


@pytest.mark.gpu
def test_gpu_render_uses_the_reference_table_semantics(fluid, oracle):
    """stale start_indices entries (the table is never cleared, src/simulation.rs:204-209) are walked exactly as the
    shader walks them: upload garbage into the table, tick once, and the render must still equal the oracle's render
    of the same (particles, table)."""
    s, ts, p = scenes.uniform_box(48, tex=(64, 64))
    rng = np.random.default_rng(3)
    inv = F.inverse_projection(s.size)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        sim.upload_start_indices(rng.integers(0, s.particle_count, size=sim.grid_w * sim.grid_h, dtype=np.uint32))
        sim.tick(ts)
        got, _, _ = F.render(sim, inv, 200, 200)
        particles, table, u = sim.download_particles(), sim.download_start_indices(), sim.uniform()
    ref = oracle.render_fluid(u, particles, table, inv, 200, 200)
    assert np.abs(got - ref).max() <= TOL

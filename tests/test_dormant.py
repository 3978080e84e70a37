"""the reference's dormant physics and scene control — surface tension (compute.wgsl:300-498,
call commented out at :92), the cursor -> world mapping that feeds the mouse force (src/main.rs:287-303) and
restart (src/renderer.rs:870-872). CPU: the oracle's restatement behaves as the shader's arithmetic implies.
GPU: the CUDA kernel is bit-identical to it on the sim's own sorted state."""
import numpy as np
import pytest

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, field as F, scenes


def _wide_scene(n_side=40, coincident=0):
    """h > 1 so that poly6_kernel_gradient(h, unit direction) is non-zero (see test_oracle_term_vanishes_for_h_le_1)."""
    s, ts, p = scenes.uniform_box(n_side, spacing=0.6, h=1.5, tex=(64, 64))
    if coincident:
        p["position"][1:1 + coincident] = p["position"][0]  # dst == 0 -> the xorshift direction (compute.wgsl:430-436)
        p["predicted_position"][1:1 + coincident] = p["predicted_position"][0]
    return s, ts, p


def test_oracle_term_vanishes_for_h_le_1(oracle):
    # the gradient kernel is handed the unit direction (compute.wgsl:440): r_len = 1 >= h -> 0 for every neighbour
    s, ts, p = scenes.dam_break(24, 48, tex=(64, 64))
    sim = oracle.OracleSim(s, p)
    sim.tick(ts)
    f = oracle.surface_tension(sim.last_uniform, sim.p, sim.table)
    assert f.shape == (s.particle_count, 2) and (f == 0).all()


def test_oracle_term_for_wide_kernel(oracle):
    s, ts, p = _wide_scene(24)
    sim = oracle.OracleSim(s, p)
    sim.tick(ts)
    ts0 = TickSettings(**{**ts.__dict__, "surface_tension_treshold": 0.0})
    sim.tick(ts0)
    f = oracle.surface_tension(sim.last_uniform, sim.p, sim.table)
    assert np.isfinite(f).all()
    mag = np.hypot(f[:, 0], f[:, 1])
    assert (mag > 0).mean() > 0.5
    # independent all-pairs float64 evaluation of the same formulas (no grid, no table, numpy broadcasting)
    u = sim.last_uniform
    pos = sim.p["predicted_position"].astype(np.float64)
    rho = sim.p["density"].astype(np.float64)
    h, PI = float(u.smoothing_radius), 3.14159265359
    off = pos[None, :, :] - pos[:, None, :]            # [i, j] = neighbour j - particle i
    sq = (off ** 2).sum(-1)
    near = sq <= float(u.sqr_radius)
    dst = np.sqrt(sq)
    assert (dst[~np.eye(len(pos), dtype=bool)] > 0).all()  # no coincident pairs here: the RNG branch is the GPU test's job
    eye = np.eye(len(pos), dtype=bool)
    d = off / np.where(eye, 1.0, dst)[..., None]
    # self (dst == 0) takes the first two xorshift32 draws of seed u32(px)*324 + frame_time*5632 (compute.wgsl:411, :430-436)
    def xorshift(x):
        x ^= (x << 13) & 0xFFFFFFFF; x ^= x >> 17; x ^= (x << 5) & 0xFFFFFFFF
        return x
    for i, px in enumerate(sim.p["predicted_position"][:, 0]):
        seed = (int(max(float(px), 0.0)) * 324 + int(u.frame_time) * 5632) & 0xFFFFFFFF
        a = xorshift(seed); b = xorshift(a)
        r = np.array([a, b], dtype=np.float64) / 4294967296.0
        d[i, i] = r / np.linalg.norm(r)
    w = float(u.particle_mass) / rho[None, :]
    grad_c = -24.0 / (PI * h ** 8) * (h * h - 1.0) ** 2  # the kernel sees the unit direction, h > 1
    lap_k = 8.0 / (PI * h ** 8) * (h * h - sq) * (3 * h * h - 4 * sq)
    lap = np.where(near, w * lap_k, 0.0).sum(1)
    n = np.where(near[..., None], (w * grad_c)[..., None] * d, 0.0).sum(1)
    n_len = np.linalg.norm(n, axis=1)
    want = (float(u.surface_tension_coefficient) * lap / (n_len + 1e-6))[:, None] * (n / n_len[:, None])
    ok = n_len > 1e-3 * np.abs(w[0] * grad_c)  # skip near-cancelled gradients: their direction is rounding noise
    assert ok.mean() > 0.9
    assert np.allclose(f[ok], want[ok], rtol=2e-3, atol=2e-3 * np.abs(want[ok]).max(axis=1, keepdims=True))
    # above a huge threshold nothing passes (compute.wgsl:304)
    sim.tick(TickSettings(**{**ts.__dict__, "surface_tension_treshold": 1e9}))
    assert (oracle.surface_tension(sim.last_uniform, sim.p, sim.table) == 0).all()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["wide", "wide-coincident", "wide-threshold", "reference-h"])
def test_gpu_surface_tension_bit_exact(fluid, oracle, case):
    if case == "reference-h":
        s, ts, p = scenes.dam_break(64, 128, tex=(64, 64))
    else:
        s, ts, p = _wide_scene(48, coincident=5 if case == "wide-coincident" else 0)
    if case != "wide-threshold":
        ts = TickSettings(**{**ts.__dict__, "surface_tension_treshold": 0.0})
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        for _ in range(2):
            sim.tick(ts)
        got = F.surface_tension(sim)
        particles, table, u = sim.download_particles(), sim.download_start_indices(), sim.uniform()
        before = particles.copy()
        assert (sim.download_particles() == before).all()  # a diagnostic: the state is untouched
    ref = oracle.surface_tension(u, particles, table)
    assert got.dtype == np.float32 and got.shape == ref.shape
    assert (got.view(np.uint32) == ref.view(np.uint32)).all(), f"{(got.view(np.uint32) != ref.view(np.uint32)).sum()} words differ"
    if case == "reference-h":
        assert (got == 0).all()
    else:
        assert (got != 0).any()


@pytest.mark.gpu
def test_gpu_surface_tension_needs_a_tick(fluid):
    with fluid.FluidSimulation.new(0, SimulationSettings(particle_count=1000, size=(8.0, 8.0), texture_size=(32, 32))) as sim:
        with pytest.raises(fluid.FluidError):
            F.surface_tension(sim)


def test_mouse_world_pos_matches_reference_mapping():
    # src/main.rs:287-303 with the orthographic projection of src/renderer.rs (inverse_projection): the window centre
    # maps to the origin; top = -size.y/2 in that projection, so the window's top-left is (-sx/2, -sy/2)
    size = (53.0, 40.0)
    inv = F.inverse_projection(size)
    win = (1280.0, 720.0)
    assert np.allclose(F.mouse_world_pos((640.0, 360.0), win, inv), (0.0, 0.0), atol=1e-6)
    assert np.allclose(F.mouse_world_pos((0.0, 0.0), win, inv), (-size[0] / 2, -size[1] / 2), rtol=1e-6)
    assert np.allclose(F.mouse_world_pos((1280.0, 720.0), win, inv), (size[0] / 2, size[1] / 2), rtol=1e-6)
    # a general matrix: same result as the f32 matrix-vector product written out
    rng = np.random.default_rng(5)
    m = rng.standard_normal(16).astype(np.float32)
    m[15] = 3.0
    cur = np.array([123.0, 456.0], np.float32)
    w = np.array(win, np.float32)
    ndc = (cur / w) * np.float32(2) - np.float32(1)
    clip = np.array([ndc[0], -ndc[1], 0, 1], np.float32)
    cols = m.reshape(4, 4)  # column-major: row k of this view = column k
    r = ((cols[0] * clip[0] + cols[1] * clip[1]) + cols[2] * clip[2]) + cols[3] * clip[3]
    want = (r[:2] / r[3]).astype(np.float32)
    got = F.mouse_world_pos(cur, w, m)
    assert (got.view(np.uint32) == want.view(np.uint32)).all()


@pytest.mark.gpu
def test_gpu_mouse_force_and_restart(fluid, oracle):
    """the mouse force itself is on the tick path (compute.wgsl:99-108): drive it through the cursor mapping, compare
    with the oracle, then restart (src/renderer.rs:870-872) and get the spawn state back."""
    s, ts, p = scenes.uniform_box(48, tex=(64, 64))
    inv = F.inverse_projection(s.size)
    mp = F.mouse_world_pos((300.0, 200.0), (800.0, 600.0), inv)
    tsm = TickSettings(**{**ts.__dict__, "mouse_pos": (float(mp[0]), float(mp[1])), "mouse_state": -1, "mouse_force_radius": 1.5})
    with fluid.FluidSimulation.new(0, s) as sim:
        spawn = sim.download_particles()
        for _ in range(3):
            sim.tick(tsm)
        got = sim.download_particles()
        ref = oracle.OracleSim(s, None)
        for _ in range(3):
            ref.tick(tsm)
        assert (got["grid"] == ref.p["grid"]).all()
        assert np.abs(got["position"] - ref.p["position"]).max() <= 1e-5
        assert np.abs(got["velocity"] - ref.p["velocity"]).max() <= 1e-4
        assert np.abs(got["velocity"]).max() > 0  # the cursor is over the fluid
        sim.restart()
        assert sim.tick_count == 0
        again = sim.download_particles()
        assert (again.view(np.uint8) == spawn.view(np.uint8)).all()

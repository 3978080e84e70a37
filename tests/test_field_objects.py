"""obstacle objects -> mask -> seeds -> JFA -> force field.
CPU part: the oracle's restatement of image_shader.wgsl against closed-form geometry. GPU part: the CUDA rasteriser
and the device-resident chain against the oracle (exact: u8 mask, u32 seeds, integer-valued float field)."""
import numpy as np
import pytest

from gpu_fluid_simulation_b200 import SimulationSettings, field as F


def _scene(rng, n, size):
    objs = []
    for i in range(n):
        pos = (rng.uniform(-size[0] / 2, size[0] / 2), rng.uniform(-size[1] / 2, size[1] / 2))
        if i % 2 == 0:
            objs.append(F.circle(pos, rng.uniform(0.5, 4.0)))
        else:
            objs.append(F.rect(pos, (rng.uniform(0.5, 8.0), rng.uniform(0.5, 5.0)), rng.uniform(-3.2, 3.2)))
    return F.objects(*objs)


# ---- oracle self-checks (CPU) ---------------------------------------------------------------------------------
def test_inverse_projection_maps_clip_to_world():
    m = F.inverse_projection((53.0, 30.0)).reshape(4, 4).T  # back to row-major
    assert np.allclose(m @ np.array([1, 1, 0, 1.0]), [26.5, -15.0, m[2, 3], 1.0])   # clip (+1,+1) = right, top = -y (renderer.rs:558-561)
    assert np.allclose(m @ np.array([-1, -1, 0, 1.0]), [-26.5, 15.0, m[2, 3], 1.0])


def test_oracle_mask_matches_closed_form_geometry(oracle):
    size, w, h = (53.0, 30.0), 424, 240
    inv = F.inverse_projection(size)
    objs = F.objects(F.circle((3.0, -2.0), 5.0), F.rect((-12.0, 4.0), (10.0, 4.0), 0.5), F.circle((100.0, 100.0), 1.0))
    mask = oracle.rasterize_objects(objs, inv, w, h)
    assert set(np.unique(mask)) == {0, 255}
    xs = ((np.arange(w) + 0.5) / w - 0.5) * size[0]
    ys = ((np.arange(h) + 0.5) / h - 0.5) * size[1]  # row 0 = clip y +1 = world y -15 (orthographic top = -sy/2)
    X, Y = np.meshgrid(xs, ys)
    circ = (X - 3.0) ** 2 + (Y + 2.0) ** 2 < 25.0
    lx, ly = X + 12.0, Y - 4.0
    c, s = np.cos(-0.5), np.sin(-0.5)
    rx, ry = lx * c - ly * s, lx * s + ly * c
    rct = (np.abs(rx) <= 5.0) & (np.abs(ry) <= 2.0)
    want = np.where(circ | rct, 0, 255)
    # fp32 vs fp64 can only disagree on pixels whose centre lies within rounding of an edge
    assert (mask != want).sum() <= 4
    assert abs((mask == 0).sum() * (size[0] / w) * (size[1] / h) - (np.pi * 25 + 40)) < 2.0
    # unknown kinds are ignored by the shader's if/else chain
    odd = objs.copy()
    odd["kind"] = 7
    assert (oracle.rasterize_objects(odd, inv, w, h) == 255).all()
    assert (oracle.rasterize_objects(F.objects(), inv, 16, 8) == 255).all()


def test_chain_on_oracle_mask_is_consistent(oracle):
    """seed rule: pixels > 128 (outside every object) are the seeds, so the field is zero outside objects and points
    from an inside pixel to the nearest outside pixel; JFA and the reference's CPU sweep agree on almost every pixel."""
    size, w, h = (53.0, 30.0), 256, 128
    mask = oracle.rasterize_objects(F.objects(F.circle((0.0, 0.0), 6.0), F.rect((12.0, 3.0), (9.0, 5.0), 0.3)),
                                    F.inverse_projection(size), w, h)
    seeds = oracle.jfa_run(oracle.jfa_seed_from_mask(mask))
    fld = oracle.jfa_to_force_field(seeds)
    assert (fld[mask == 255] == 0).all()
    inside = mask == 0
    assert (np.abs(fld[inside]).sum(axis=1) > 0).all()
    sweep = oracle.generate_smooth_gradient_field(mask)
    d_jfa = np.hypot(fld[..., 0], fld[..., 1])
    d_swp = np.hypot(sweep[..., 0], sweep[..., 1])
    assert np.mean(np.abs(d_jfa - d_swp) > 1e-6) < 0.02  # both are near-exact nearest-seed transforms


# ---- CUDA vs oracle ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("w,h,n", [(64, 32, 0), (257, 131, 3), (1024, 1024, 9), (960, 540, 200)])
def test_gpu_mask_equals_oracle(fluid, oracle, w, h, n):
    rng = np.random.default_rng(w + n)
    size = (53.0, 30.0)
    objs = _scene(rng, n, size)
    inv = F.inverse_projection(size)
    assert np.array_equal(F.raster_objects(objs, inv, w, h), oracle.rasterize_objects(objs, inv, w, h))


@pytest.mark.gpu
def test_gpu_mask_with_general_matrix(fluid, oracle):
    """a non-trivial inv_proj (rotation + perspective-like w row): every matrix term of image_shader.wgsl:44-45 is used."""
    rng = np.random.default_rng(5)
    m = np.eye(4) + 0.2 * rng.standard_normal((4, 4))
    m[3] = [0.05, -0.03, 0.0, 1.0]
    inv = np.ascontiguousarray(m.T.astype(np.float32)).reshape(16)
    objs = _scene(rng, 12, (2.0, 2.0))
    for o in objs:
        o["position"] *= 0.5
    assert np.array_equal(F.raster_objects(objs, inv, 300, 200), oracle.rasterize_objects(objs, inv, 300, 200))


@pytest.mark.gpu
@pytest.mark.parametrize("tex,n", [((256, 128), 5), ((1024, 1024), 7), ((128, 128), 0)])
def test_gpu_chain_equals_oracle_and_feeds_the_step(fluid, oracle, tex, n):
    rng = np.random.default_rng(tex[0] + n)
    size = (53.0, 30.0)
    s = SimulationSettings(particle_count=2000, size=size, texture_size=tex)
    objs = _scene(rng, n, size)
    inv = F.inverse_projection(size)
    ref_mask = oracle.rasterize_objects(objs, inv, tex[0], tex[1])
    ref_seeds = oracle.jfa_run(oracle.jfa_seed_from_mask(ref_mask))
    ref_field = oracle.jfa_to_force_field(ref_seeds)
    with fluid.FluidSimulation.new(0, s) as sim:
        mask, seeds, fld, ms = F.field_from_objects(sim, objs, inv, want_mask=True, want_seeds=True, want_field=True, timed=True)
        assert np.array_equal(mask, ref_mask)
        assert np.array_equal(seeds, ref_seeds)
        assert np.array_equal(fld, ref_field)
        assert ms > 0
        # asynchronous form (no outputs), then read back what the step will consume
        F.field_from_objects(sim, objs, inv)
        assert np.array_equal(sim.download_force_field().reshape(tex[1], tex[0], 2), ref_field)
        # the mask entry point (reference `Image`) gives the same field
        seeds2, fld2, _ = F.field_from_mask(sim, ref_mask, want_seeds=True, want_field=True)
        assert np.array_equal(seeds2, ref_seeds) and np.array_equal(fld2, ref_field)
        # and one tick with that field equals the oracle's tick with the oracle's field (push-out path, compute.wgsl:127-140)
        from gpu_fluid_simulation_b200 import TickSettings
        ts = TickSettings()
        p0 = sim.download_particles()
        osim = oracle.OracleSim(s, p0)
        osim.texture = ref_field.copy()
        sim.tick(ts)
        osim.tick(ts)
        got = sim.download_particles()
        for f in ("position", "velocity", "density", "predicted_position", "grid"):
            assert np.array_equal(got[f], osim.p[f]), f

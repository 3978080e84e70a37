"""GPU parity for the jump-flood field: exact equality with the oracle (criterion 3)."""
import os

import numpy as np
import pytest

from gpu_fluid_simulation_b200 import scenes

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _random_seeds(rng, h, w, frac):
    mask = np.where(rng.random((h, w)) < frac, 255, 0).astype(np.uint8)
    return mask


@pytest.mark.parametrize("h,w,frac", [(1, 1, 1.0), (7, 5, 0.2), (50, 70, 0.03), (64, 64, 0.001), (129, 257, 0.01),
                                      (256, 256, 0.0002), (300, 200, 0.5), (1024, 1024, 0.0001)])
def test_full_run_matches_oracle(fluid, oracle, h, w, frac):
    rng = np.random.default_rng(h * 1000 + w)
    mask = _random_seeds(rng, h, w, frac)
    seeds = oracle.jfa_seed_from_mask(mask)
    assert np.array_equal(fluid.jfa.seed_from_mask(mask), seeds)
    ref = oracle.jfa_run(seeds)
    got = fluid.jfa.jfa_run(seeds)
    assert np.array_equal(got, ref)
    assert np.array_equal(fluid.jfa.to_force_field(got), oracle.jfa_to_force_field(ref))


@pytest.mark.parametrize("jump", [1, 2, 3, 4, 8, 64, 128])
def test_single_pass_matches_oracle(fluid, oracle, jump):
    rng = np.random.default_rng(jump)
    for h, w in ((96, 160), (97, 161)):
        seeds = oracle.jfa_seed_from_mask(_random_seeds(rng, h, w, 0.02))
        assert np.array_equal(fluid.jfa.jfa_pass(seeds, jump), oracle.jfa_pass(seeds, jump))


def test_no_seed_and_border_fallback(fluid, oracle):
    mask = np.zeros((33, 47), np.uint8)  # no white pixel: border seeds (src/main.rs:425-438)
    seeds = oracle.jfa_seed_from_mask(mask)
    assert np.array_equal(fluid.jfa.seed_from_mask(mask), seeds)
    assert np.array_equal(fluid.jfa.jfa_run(seeds), oracle.jfa_run(seeds))
    inv = np.full((20, 20, 2), 32767, np.uint32)  # nothing valid anywhere: stays invalid, field zero
    out = fluid.jfa.jfa_run(inv)
    assert np.array_equal(out, oracle.jfa_run(inv))
    assert not fluid.jfa.to_force_field(out).any()


def test_golden_jfa(fluid):
    with np.load(os.path.join(GOLDEN, "jfa_70x50.npz")) as z:
        assert np.array_equal(fluid.jfa.seed_from_mask(z["mask"]), z["seeds"])
        assert np.array_equal(fluid.jfa.jfa_pass(z["seeds"], 8), z["pass8"])
        out = fluid.jfa.jfa_run(z["seeds"])
        assert np.array_equal(out, z["out"])
        assert np.array_equal(fluid.jfa.to_force_field(out), z["field"])


def test_seeds_from_particles_and_sim_pipeline(fluid, oracle):
    s, ts, p = scenes.dam_break(128, 512, tex=(256, 256))
    ref_seeds = oracle.jfa_seed_from_particles(p, s.size, 256, 256)
    assert np.array_equal(fluid.jfa.seed_from_particles(p, s.size, 256, 256), ref_seeds)
    ref = oracle.jfa_run(ref_seeds)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        seeds, fld, ms = sim.jfa_from_particles(256, 256, into_sim=True)
        assert np.array_equal(seeds, ref)
        assert np.array_equal(fld, oracle.jfa_to_force_field(ref))
        assert np.array_equal(sim.download_force_field(), fld)
        assert ms > 0


@pytest.mark.parametrize("size,n_side", [(4096, 1000)])
def test_config4_jfa_4096_from_particles(fluid, oracle, size, n_side):
    """BASELINE.json config 4 (4096^2): dense seeds (uniform box) and long-range field (dam-break)."""
    for s, ts, p in (scenes.uniform_box(n_side), scenes.dam_break(n_side // 2, n_side * 2)):
        ref = oracle.jfa_run(oracle.jfa_seed_from_particles(p, s.size, size, size))
        with fluid.FluidSimulation.new(0, s) as sim:
            sim.upload_particles(p)
            seeds, fld, ms = sim.jfa_from_particles(size, size)
        assert np.array_equal(seeds, ref)
        assert np.array_equal(fld, oracle.jfa_to_force_field(ref))


@pytest.mark.parametrize("scene", ["4m-d", "4m-u"])
def test_config4_jfa_8192_from_4m_particles(fluid, oracle, scene):
    """BASELINE.json config 4 as stated: the 8192^2 field (13 passes, shaders/jump_flood.wgsl:17-58) from 4M particles —
    the long-range field of the 4M dam break and the dense seeds of the 4M uniform box — seed map and derived field
    exactly equal to the oracle's (a few seconds of CPU per scene)."""
    s, ts, p = scenes.dam_break(1000, 4000) if scene == "4m-d" else scenes.uniform_box(2000)
    assert p.shape[0] == 4_000_000
    size = 8192
    ref_seeds = oracle.jfa_seed_from_particles(p, s.size, size, size)
    ref = oracle.jfa_run(ref_seeds)
    with fluid.FluidSimulation.new(0, s) as sim:
        sim.upload_particles(p)
        seeds, fld, ms = sim.jfa_from_particles(size, size)
    assert np.array_equal(seeds, ref), f"{int((seeds != ref).any(axis=2).sum())} pixels of the 8192^2 seed map differ"
    assert np.array_equal(fld, oracle.jfa_to_force_field(ref))
    # the host-buffer entry point on the same seeds (the reference-format rg32uint image in, rg32uint out)
    assert np.array_equal(fluid.jfa.jfa_run(ref_seeds), ref)
    assert ms > 0


def test_non_canonical_seed_images_take_the_general_path(fluid, oracle):
    """caller-supplied seed values the shader would skip as neighbours (coordinate > width) but that are not the
    INVALID marker: the fast kernels' no-skip argument does not hold, the library must detect them while packing and
    fall back to the general kernels. Also partially-invalid pairs and sx == width (admitted by the shader's `>`)."""
    rng = np.random.default_rng(11)
    for h, w in ((64, 64), (96, 128), (128, 96)):
        seeds = oracle.jfa_seed_from_mask(_random_seeds(rng, h, w, 0.02))
        odd = rng.random((h, w)) < 0.01
        seeds[odd, 0] = rng.integers(w + 1, 32768, size=int(odd.sum()))       # x beyond the width (<= 32767: the shader's i32 distance must not overflow)
        half = rng.random((h, w)) < 0.005
        seeds[half, 1] = 32767                                                 # only one channel invalid
        edge = rng.random((h, w)) < 0.005
        seeds[edge] = (w, min(h - 1, w))                                       # sx == width: not skipped
        assert np.array_equal(fluid.jfa.jfa_run(seeds), oracle.jfa_run(seeds))
        for jump in (1, 2, 4, 16):
            assert np.array_equal(fluid.jfa.jfa_pass(seeds, jump), oracle.jfa_pass(seeds, jump))


@pytest.mark.parametrize("h,w", [(512, 512), (257, 256), (255, 256), (100, 4), (8, 8)])
def test_canonical_fast_path_every_jump(fluid, oracle, h, w):
    """images the window / no-skip kernels are used for (w % 4 == 0, h <= w + 1), pass by pass incl. jumps 1 and 2 at
    the image borders, dense and sparse seeds."""
    rng = np.random.default_rng(h * 7 + w)
    for frac in (0.3, 0.002):
        seeds = oracle.jfa_seed_from_mask(_random_seeds(rng, h, w, frac))
        for jump in (1, 2, 4, 8, 64):
            assert np.array_equal(fluid.jfa.jfa_pass(seeds, jump), oracle.jfa_pass(seeds, jump)), (jump, frac)
        assert np.array_equal(fluid.jfa.jfa_run(seeds), oracle.jfa_run(seeds))

"""fluid_jfa_run_static — the jump flood with its scratch in a file-scope `static __device__` array (no cudaMalloc in the
call; include/fluid_b200.h) — against the existing entry point fluid_jfa_run, through the C ABI, on calls in a row.

The entry point has no counterpart in the reference; the result it must
reproduce is fluid_jfa_run's, which tests/test_gpu_jfa.py pins against the oracle (shaders/jump_flood.wgsl:17-65). JFA
is integer work: the comparison is bit for bit, as everywhere else in the JFA tests.

Collected last (the file name sorts after every other test file).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1 << 20  # FLUID_JFA_STATIC_MAX_PIXELS
INVALID = 32767


def _seeds(w, h, n_seeds, rng):
    img = np.full((h, w, 2), INVALID, dtype=np.uint32)
    ys = rng.integers(0, h, n_seeds)
    xs = rng.integers(0, w, n_seeds)
    img[ys, xs, 0] = xs
    img[ys, xs, 1] = ys
    return img


def _run_existing(lib, _lib, seeds):
    h, w = seeds.shape[:2]
    out = np.empty_like(seeds)
    _lib.check(lib.fluid_jfa_run(_lib.ptr(seeds), w, h, _lib.ptr(out), 0))
    return out


def _run_static_host(lib, _lib, seeds):
    h, w = seeds.shape[:2]
    out = np.zeros_like(seeds)
    _lib.check(lib.fluid_jfa_run_static(_lib.ptr(seeds), w, h, _lib.ptr(out), 0, None, 1))
    return out


def test_static_scratch_matches_fluid_jfa_run_two_calls_in_a_row(fluid):
    """host buffers, like fluid_jfa_run's: a large image, then a smaller ragged one that leaves the first call's
    values in the rest of the shared array, then the first again."""
    from gpu_fluid_simulation_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(41)
    first = _seeds(1024, 1024, 300, rng)      # exactly FLUID_JFA_STATIC_MAX_PIXELS
    second = _seeds(333, 517, 40, rng)        # w % 4 != 0: the scalar passes
    want_first, want_second = _run_existing(lib, _lib, first), _run_existing(lib, _lib, second)
    assert np.array_equal(_run_static_host(lib, _lib, first), want_first)
    assert np.array_equal(_run_static_host(lib, _lib, second), want_second)
    assert np.array_equal(_run_static_host(lib, _lib, first), want_first)


def test_static_scratch_non_canonical_and_empty_images(fluid):
    """seed images the shader treats specially: no seed at all, and out-of-range markers other than INVALID
    (jump_flood.wgsl:48 skips sx > width or sy > width)."""
    from gpu_fluid_simulation_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(42)
    empty = np.full((64, 96, 2), INVALID, dtype=np.uint32)
    odd = _seeds(200, 120, 25, rng)
    odd[5, 7] = (40000, 3)    # skipped by the shader: x > width
    odd[9, 11] = (200, 100)   # x == width is NOT skipped (`>`), although it lies outside the image
    for img in (empty, odd, empty):
        assert np.array_equal(_run_static_host(lib, _lib, img), _run_existing(lib, _lib, img))


def test_static_scratch_device_pointers_enqueue_only(fluid):
    """device-resident input and output (the caller that cannot allocate already owns them): the call only enqueues on
    the caller's stream; two calls back to back on that stream, then one synchronisation."""
    torch = pytest.importorskip("torch")
    from gpu_fluid_simulation_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(43)
    imgs = [_seeds(512, 384, 60, rng), _seeds(640, 640, 10, rng)]
    want = [_run_existing(lib, _lib, im) for im in imgs]
    dev = torch.device("cuda:0")
    stream = torch.cuda.Stream(device=dev)
    d_in = [torch.from_numpy(im.view(np.int32)).to(dev) for im in imgs]
    d_out = [torch.zeros_like(t) for t in d_in]
    torch.cuda.synchronize(dev)
    for t_in, t_out, im in zip(d_in, d_out, imgs):
        h, w = im.shape[:2]
        _lib.check(lib.fluid_jfa_run_static(C.c_void_p(t_in.data_ptr()), w, h, C.c_void_p(t_out.data_ptr()), 0,
                                            C.c_void_p(stream.cuda_stream), 0))
    stream.synchronize()
    for t_out, w_ in zip(d_out, want):
        assert np.array_equal(t_out.cpu().numpy().view(np.uint32), w_)


def test_static_scratch_rejects_images_beyond_the_maximum(fluid):
    from gpu_fluid_simulation_b200 import _lib
    lib = _lib.load()
    w, h = 1024, 1025
    assert w * h > MAX_PIXELS
    seeds = np.full((h, w, 2), INVALID, dtype=np.uint32)
    out = np.zeros_like(seeds)
    rc = lib.fluid_jfa_run_static(_lib.ptr(seeds), w, h, _lib.ptr(out), 0, None, 1)
    assert rc == -4  # FLUID_ERR_UNSUPPORTED
    msg = lib.fluid_last_error().decode()
    assert "FLUID_JFA_STATIC_MAX_PIXELS" in msg and "1048576" in msg
    assert not out.any()  # nothing was written
    # and the entry point still works afterwards
    ok = _seeds(128, 128, 9, np.random.default_rng(44))
    assert np.array_equal(_run_static_host(lib, _lib, ok), _run_existing(lib, _lib, ok))

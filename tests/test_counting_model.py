"""CPU check of the counting-sort logic of kernels_count.cuh (run-aggregated slot assignment under arbitrary atomic
orderings, then the in-cell rank by source index): the numpy model in scripts/model_counting_sort.py must give the stable
sort for sorted-ish, random and single-cell key arrays. The kernels themselves are tested on the GPU
(tests/test_gpu_sort_paths.py); this pins the algorithm they implement."""
import importlib.util
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _model():
    spec = importlib.util.spec_from_file_location("model_counting_sort", os.path.join(ROOT, "scripts", "model_counting_sort.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_model_is_the_stable_sort():
    _model().main()


def test_model_ragged_tail_and_out_of_grid_keys():
    m = _model()
    rng = np.random.default_rng(3)
    for n in (1, 31, 32, 33, 95, 129):
        keys = rng.integers(0, 12, n).astype(np.int64)
        keys[rng.random(n) < 0.1] = 40  # beyond the grid: clamped into the last cell for the tables, key kept
        sk, perm, lb = m.model(keys, 10, rng)
        kc = np.minimum(keys, 9)
        want = np.argsort(kc, kind="stable")
        assert np.array_equal(perm, want)
        assert np.array_equal(lb, np.concatenate([[0], np.cumsum(np.bincount(kc, minlength=10))]))

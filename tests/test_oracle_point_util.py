"""CPU: the libPointUtil restatement (oracle/point_util_oracle.py) against plain numpy on small cases — the C loops
(cross radius search, k-NN) vs argsort / boolean masks, and the structural properties the call sites rely on."""
from __future__ import annotations

import numpy as np

from oracle import point_util_oracle as O


def test_knn_matches_stable_argsort():
    rng = np.random.default_rng(0)
    s = rng.uniform(0, 1, (700, 3)).astype(np.float32)
    s[10:20] = s[0:10]  # ties
    q = rng.uniform(0, 1, (150, 3)).astype(np.float32)
    idxs, d2 = O.knn(s, q, 8)
    dx, dy, dz = (s[None, :, k] - q[:, None, k] for k in range(3))
    ref = (dz.astype(np.float64) * dz + (dx.astype(np.float64) * dx + (dy * dy).astype(np.float64)).astype(np.float32)).astype(np.float32)
    order = np.argsort(ref, axis=1, kind="stable")[:, :8]
    assert np.array_equal(idxs, order)
    assert np.array_equal(d2, np.take_along_axis(ref, order, 1))


def test_cross_search_matches_mask():
    rng = np.random.default_rng(1)
    s = rng.uniform(0, 1, (400, 3)).astype(np.float32)
    q = rng.uniform(0, 1, (90, 3)).astype(np.float32)
    idxs, lens, begs, cens = O.cross_search(s, q, 0.04)
    d2 = ((s[None].astype(np.float64) - q[:, None]) ** 2).sum(-1)
    for i in range(90):
        row = idxs[begs[i]:begs[i] + lens[i]]
        sure = np.nonzero(d2[i] < 0.04 * (1 - 1e-5))[0]
        maybe = np.nonzero(d2[i] < 0.04 * (1 + 1e-5))[0]
        assert set(sure) <= set(row) <= set(maybe) and np.all(np.diff(row) > 0)
    assert np.array_equal(cens, np.repeat(np.arange(90), lens))


def test_grid_downsample_one_lowest_point_per_voxel():
    rng = np.random.default_rng(2)
    x = rng.uniform(0, 2, (5000, 3)).astype(np.float32)
    idx = O.grid_downsample(x, 0.25)
    v = np.floor((x - x.min(0)) / np.float32(0.25)).astype(np.int64)
    keys = v[:, 0] * 10000 + v[:, 1] * 100 + v[:, 2]
    assert len(np.unique(keys)) == len(idx) and np.all(np.diff(keys[idx]) > 0)
    first = {}
    for i, k in enumerate(keys):
        first.setdefault(k, i)
    assert sorted(first.values()) == sorted(idx.tolist())


def test_covars_and_interpolation_small():
    x = np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0], [0, 1, 0]], np.float32)
    c = O.compute_covars(x, np.array([0, 1, 2], np.int32), np.array([3], np.int32), np.array([0], np.int32))
    assert np.allclose(c[0], [1, 0, 0, 0, 0, 0, 0, 0, 0])  # points on the x axis: all variance in xx
    pr = np.eye(4, dtype=np.float32)
    q = np.array([[0.1, 0, 0]], np.float32)
    out = O.interpolate_probs(x, q, pr, np.array([0, 1], np.int32), np.array([2], np.int32), np.array([0], np.int32), 1.0)
    w = np.exp(-np.array([0.01, 0.81]))
    assert np.allclose(out[0, :2], w / w.sum()) and not out[0, 2:].any()


def test_block_origins_as_the_reference():
    assert O.block_origins(2.0, 3.0, 1.5) == [0.0]
    assert O.block_origins(7.0, 3.0, 1.5) == [0.0, 1.5, 3.0, 4.0]
    assert O.block_origins(6.0, 3.0, 1.5) == [0.0, 1.5, 3.0]

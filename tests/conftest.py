"""pytest configuration: the `gpu` marker, repo root on sys.path, shared synthetic generators."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


# ---- synthetic inputs (seeded; the reference's tests draw the same shapes unseeded) -----------------
def random_csr(rng, pn, max_len, min_len=0):
    """lens in [min_len,max_len) incl. empty rows, neighbours without replacement
    (tf_ops/test/test_neighbor_scatter.py:171-185)."""
    lens = np.minimum(rng.integers(min_len, max_len, pn), pn).astype(np.int32)
    rows = [rng.choice(pn, int(n), replace=False) for n in lens if n > 0]
    nidxs = (np.concatenate(rows) if rows else np.zeros(0)).astype(np.int32)
    bgs = (np.cumsum(lens, dtype=np.int64) - lens).astype(np.int32)
    cens = np.repeat(np.arange(pn, dtype=np.int32), lens)
    return nidxs, lens, bgs, cens


def room_block(rng, pn, block=3.0, height=2.7):
    """S3DIS-like block: points on floor/ceiling/4 walls + 2 % volume noise, xy in [-block/2, block/2],
    z in [0, height] (pre-normalised as s3dis_util.py:118-120 does)."""
    n_noise = max(1, pn // 50)
    n_surf = pn - n_noise
    face = rng.integers(0, 6, n_surf)
    u = rng.uniform(-block / 2, block / 2, n_surf)
    v = rng.uniform(-block / 2, block / 2, n_surf)
    w = rng.uniform(0, height, n_surf)
    pts = np.empty((n_surf, 3), np.float64)
    for f in range(6):
        m = face == f
        if f == 0:
            pts[m] = np.stack([u[m], v[m], np.zeros(m.sum())], 1)
        elif f == 1:
            pts[m] = np.stack([u[m], v[m], np.full(m.sum(), height)], 1)
        elif f == 2:
            pts[m] = np.stack([np.full(m.sum(), -block / 2), v[m], w[m]], 1)
        elif f == 3:
            pts[m] = np.stack([np.full(m.sum(), block / 2), v[m], w[m]], 1)
        elif f == 4:
            pts[m] = np.stack([u[m], np.full(m.sum(), -block / 2), w[m]], 1)
        else:
            pts[m] = np.stack([u[m], np.full(m.sum(), block / 2), w[m]], 1)
    noise = np.stack([rng.uniform(-block / 2, block / 2, n_noise), rng.uniform(-block / 2, block / 2, n_noise),
                      rng.uniform(0, height, n_noise)], 1)
    pts = np.concatenate([pts, noise])
    pts += rng.normal(0, 0.005, pts.shape)
    pts[:, :2] = np.clip(pts[:, :2], -block / 2, block / 2)
    pts[:, 2] = np.clip(pts[:, 2], 0, None)
    return pts[rng.permutation(pn)].astype(np.float32)


@pytest.fixture
def rng():
    return np.random.default_rng(1234)

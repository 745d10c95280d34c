"""CPU: the numpy restatement of the pooling composites (oracle/cpu_pooling.py) satisfies the reference's
own structural checks (PoolingIndex/test_permutation.py:55-88, test_classes_permutation.py:47-75,
model_pooling.py:774-799) at every level, and its N-level builder agrees with a literal line-by-line
transcription of the two-level script (graph_pooling_layer.py:110-146) written on the raw oracle calls."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import room_block
from oracle import cpu_pooling as cp
from oracle import pcs_oracle as orc


def _block(seed, pn):
    rng = np.random.default_rng(seed)
    return (rng, room_block(rng, pn), rng.uniform(-1, 1, (pn, 5)).astype(np.float32),
            rng.integers(0, 13, pn).astype(np.int64))


def _literal_two_layers(x, f, lab, v1, v2, bs):
    lab = lab.astype(np.int32)[:, None]
    o2p1, p2o1, vl1, vb1, vc1 = orc.permutation_info(orc.voxel_index(x, v1, bs))
    pts1, f, lab = orc.row_gather(x, o2p1), orc.row_gather(f, o2p1), orc.row_gather(lab, o2p1)
    pts2 = cp.centroid(pts1, vl1, vb1)
    d1 = orc.diff_xyz(pts1, pts2, vc1)
    o2p2, p2o2, vl2, vb2, vc2 = orc.permutation_info(orc.voxel_index(pts2, v2, bs))
    pts2 = orc.row_gather(pts2, o2p2)
    pts3 = cp.centroid(pts2, vl2, vb2)
    d2 = orc.diff_xyz(pts2, pts3, vc2)
    ro2p, rp2o, vl1, vb1, vc1 = orc.repermutation_info(o2p2, vl1, vb1, x.shape[0])
    pts1, d1, f, lab = (orc.row_gather(a, ro2p) for a in (pts1, d1, f, lab))
    return [pts1, pts2, pts3], [d1, d2], f, lab[:, 0].astype(np.int64), [vl1, vl2], [vb1, vb2], [vc1, vc2]


def _flat(o):
    for x in o:
        if isinstance(x, (list, tuple)):
            yield from _flat(x)
        else:
            yield x


def test_hierarchy_builder_equals_literal_script():
    _, x, f, lab = _block(1, 5000)
    a = cp.points_pooling_two_layers(x, f, lab, 0.15, 0.45, 3.0)
    b = _literal_two_layers(x, f, lab, 0.15, 0.45, 3.0)
    for u, v in zip(_flat(a), _flat(b)):
        assert u.dtype == v.dtype and np.array_equal(u, v)


@pytest.mark.parametrize("voxels", [(0.15, 0.45), (0.15, 0.45, 0.9), (0.1, 0.3, 1.5)])
def test_structural_invariants_every_level(voxels):
    _, x, f, lab = _block(2, 6000)
    fn = cp.points_pooling_two_layers if len(voxels) == 2 else cp.points_pooling_three_layers
    xs, ds, pf, plab, vl, vb, vc = fn(x, f, lab, *voxels, 3.0)
    assert len(xs) == len(voxels) + 1
    for k in range(len(voxels)):
        cp.check_vidxs(vl[k], vb[k], vc[k], xs[k].shape[0])
        cp.check_dxyzs(xs[k], xs[k + 1], ds[k], vc[k])
        assert np.array_equal(cp.centroid(xs[k], vl[k], vb[k]), xs[k + 1])
    # the finest level is a permutation of the input that carried features and labels with it
    order = np.lexsort(xs[0].T[::-1]); order0 = np.lexsort(x.T[::-1])
    assert np.array_equal(xs[0][order], x[order0])
    assert np.array_equal(pf[order], f[order0]) and np.array_equal(plab[order], lab[order0])


def test_single_level_composites():
    rng, x, f, lab = _block(3, 3000)
    ppts, cpts, dpts, pf, plab, vl, vb, vc = cp.points_pooling(x, f, lab, 0.2, 3.0)
    cp.check_vidxs(vl, vb, vc, 3000); cp.check_dxyzs(ppts, cpts, dpts, vc)
    ctx = rng.integers(0, 3000, 100)
    out = cp.context_points_pooling(x, f, ctx, 0.2, 3.0)
    assert np.array_equal(out[0][out[-1]], x[ctx])
    cen, cf = cp.average_downsample(x, f, 0.2, 1.5)
    assert cen.shape[0] == cf.shape[0] == len(vl) and cf.shape[1] == 5
    classes = rng.integers(0, 4, 3000)
    o = cp.class_pooling(x, f, classes, lab, 0.3, 3.0)
    cp.check_vidxs(o[5], o[6], o[7], 3000)
    assert len(o[5]) >= len(cp.points_pooling(x, f, lab, 0.3, 3.0)[5])  # class split only adds voxels


def test_centroid_is_sequential_fp32():
    rng = np.random.default_rng(0)
    lens = rng.integers(1, 40, 200).astype(np.int32)
    begs = orc.bgs_from_lens(lens)
    pts = rng.uniform(-1, 1, (int(lens.sum()), 3)).astype(np.float32)
    want = np.zeros((200, 3), np.float32)
    for i in range(200):
        acc = np.zeros(3, np.float32)
        for j in range(lens[i]):
            acc = acc + pts[begs[i] + j]
        want[i] = acc / np.float32(lens[i])
    assert np.array_equal(cp.centroid(pts, lens, begs), want)

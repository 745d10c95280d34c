"""GPU parity at BASELINE size (configs[2]): the 64-block S3DIS batch that bench.py trains on
(`s3dis_batch(seed, 64)`, ~515 k points, blocks of 6-10 k) goes through the block-diagonal
`points_pooling_two_layers(block_ids=...)` and the tower's nine batched searches, and every block's slice is
compared, bit for bit, with the REFERENCE's own kernels (oracle/_ref/libref.so) run one block per call,
as the reference runs them (train_graph_pool.py:95-108: one block per tower).

This is the batched path's parity at the size that is benchmarked; the small-input oracle tests
(tests/test_gpu_pooling_ops.py, test_gpu_network.py) cover the edge cases."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import libref
from pointcloudsegmentation_b200.network import _STAGES
from pointcloudsegmentation_b200.synthetic import s3dis_batch

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not libref.SO.exists(), reason="oracle/_ref/libref.so not built")]

NB = 64
f32 = np.float32


def _ref_two_layers(x, v1, v2, bs):
    """points_pooling_two_layers' index pipeline on ONE block through the reference's kernels
    (graph_pooling_layer.py:110-146); centroid = libref sum / count."""
    mn = float(-f32(bs) / f32(2.0) - f32(1e-3)); mz = float(f32(0.0) - f32(1e-3))
    o2p1, p2o1, vl1, vb1, vc1 = libref.permutation_info(libref.voxel_index(x, float(f32(v1)), mn, mn, mz))
    pts1 = libref.permutate_feature(x, o2p1)
    pts2 = libref.sum_feat_gather(pts1, vl1, vb1) / vl1.float().unsqueeze(1)
    o2p2, p2o2, vl2, vb2, vc2 = libref.permutation_info(libref.voxel_index(pts2, float(f32(v2)), mn, mn, mz))
    pts2 = libref.permutate_feature(pts2, o2p2)
    pts3 = libref.sum_feat_gather(pts2, vl2, vb2) / vl2.float().unsqueeze(1)
    ro2p, rp2o, vl1, vb1, vc1 = libref.repermutation_info(o2p2, vl1, vb1, x.shape[0])
    pts1 = libref.permutate_feature(pts1, ro2p)
    return [pts1, pts2, pts3], [vl1, vl2], [vb1, vb2], [vc1, vc2]


@pytest.fixture(scope="module")
def batch():
    from pointcloudsegmentation_b200 import graph_layers as G

    d = s3dis_batch(1000, NB)  # bench.py's rank-0 batch 0
    x, f, l, b = (torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids"))
    xs, dxs, pf, lab, vl, vb, vc, bids = G.points_pooling_two_layers(x, f, l, 0.15, 0.45, 3.0, block_ids=b)
    bbegs = [G.block_begs_from_ids(t, NB) for t in bids]
    return d, x, xs, vl, vb, vc, [bb.cpu().numpy() for bb in bbegs], bbegs


def test_batched_pooling_equals_libref_per_block(batch):
    d, x, xs, vl, vb, vc, hb, _ = batch
    assert d["xyzs"].shape[0] > 400_000
    bid = d["block_ids"]
    starts = np.concatenate([[0], np.cumsum(np.bincount(bid, minlength=NB))])
    for k in range(NB):
        r_xs, r_vl, r_vb, r_vc = _ref_two_layers(x[starts[k]:starts[k + 1]].contiguous(), 0.15, 0.45, 3.0)
        for lev in range(3):
            a, e = hb[lev][k], hb[lev][k + 1]
            assert torch.equal(xs[lev][a:e], r_xs[lev]), (k, lev)
        for lev in range(2):
            a, e = hb[lev][k], hb[lev][k + 1]
            a2, e2 = hb[lev + 1][k], hb[lev + 1][k + 1]
            assert torch.equal(vl[lev][a2:e2], r_vl[lev]), (k, lev)
            assert torch.equal(vb[lev][a2:e2] - int(a), r_vb[lev]), (k, lev)
            assert torch.equal(vc[lev][a:e] - int(a2), r_vc[lev]), (k, lev)


def test_nine_batched_searches_equal_libref_per_block(batch):
    from pointcloudsegmentation_b200 import graph_layers as G

    _, _, xs, _, _, _, hb, bbegs = batch
    for level, rmin, rmax, _, _ in _STAGES:
        bb = bbegs[level]
        mb = int((hb[level][1:] - hb[level][:-1]).max())
        if rmin is None:
            idxs, lens, begs, cens = G.search_neighborhood(xs[level], rmax, block_begs=bb, max_block_pn=mb)
        else:
            idxs, lens, begs, cens = G.search_neighborhood_range(xs[level], rmin, rmax, block_begs=bb, max_block_pn=mb)
        assert int(begs[-1] + lens[-1]) == idxs.shape[0]
        for k in range(NB):
            a, e = int(hb[level][k]), int(hb[level][k + 1])
            one = xs[level][a:e].contiguous()
            if rmin is None:
                r = libref.search(one, float(f32(rmax * rmax)))
            else:
                r = libref.search(one, float(f32(rmax * rmax)), float(f32(rmin * rmin)))
            ea = int(begs[a]); ee = ea + int(r[0].shape[0])
            assert torch.equal(lens[a:e], r[1]), (level, rmax, k)
            assert torch.equal(begs[a:e] - ea, r[2]), (level, rmax, k)
            assert torch.equal(idxs[ea:ee] - a, r[0]), (level, rmax, k)
            assert torch.equal(cens[ea:ee] - a, r[3]), (level, rmax, k)

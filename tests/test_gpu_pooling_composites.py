"""GPU parity, row a21 of SURVEY.md §8: every pooling composite of tf_ops/graph_pooling_layer.py:39-352
(`points_pooling`, `average_downsample`, `context_points_pooling`, `class_pooling`,
`points_pooling_two_layers`, `context_points_pooling_two_layers`, `points_pooling_three_layers`,
`voxel_downsample`) against the numpy restatement in oracle/cpu_pooling.py.

Bar: EVERYTHING bit-exact — indices, labels, and the coordinates too: the centroid is a sequential fp32
sum in edge order divided in fp32 in the reference kernel (TFNeighborKernelNew.cu:19-25), in the oracle and in
`k_segment_sum`, so no tolerance is needed (a 1-ulp centroid difference could move a level-2 voxel index).
Plus the reference's own structural checks at every level (PoolingIndex/test_permutation.py:55-88,
model_pooling.py:774-799)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import room_block
from oracle import cpu_pooling as cp

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def flat(o):
    for x in o:
        if isinstance(x, (list, tuple)):
            yield from flat(x)
        else:
            yield x


def same_outputs(got, want):
    got, want = list(flat(got)), list(flat(want))
    assert len(got) == len(want)
    for k, (g, w) in enumerate(zip(got, want)):
        g = g.detach().cpu().numpy()
        assert g.shape == w.shape, (k, g.shape, w.shape)
        assert g.dtype == w.dtype, (k, g.dtype, w.dtype)
        assert np.array_equal(g, w), (k, int((g != w).sum()), g.size)


def block(seed, pn, nfeat=6, ncls=13):
    rng = np.random.default_rng(seed)
    xyz = room_block(rng, pn)
    feats = rng.uniform(-1, 1, (pn, nfeat)).astype(np.float32)
    labels = rng.integers(0, ncls, pn).astype(np.int64)
    return rng, xyz, feats, labels


@pytest.fixture(scope="module")
def G():
    from pointcloudsegmentation_b200 import graph_layers
    return graph_layers


@pytest.mark.parametrize("pn,voxel", [(6000, 0.2), (1, 0.2), (2500, 0.45)])
def test_points_pooling(G, pn, voxel):
    _, xyz, feats, labels = block(pn, pn)
    got = G.points_pooling(dev(xyz), dev(feats), dev(labels), voxel, 3.0)
    want = cp.points_pooling(xyz, feats, labels, voxel, 3.0)
    same_outputs(got, want)
    ppts, cpts, dpts, _, _, vl, vb, vc = want
    cp.check_vidxs(vl, vb, vc, pn)
    cp.check_dxyzs(ppts, cpts, dpts, vc)


def test_voxel_downsample(G):
    _, xyz, feats, labels = block(3, 4000)
    got = G.voxel_downsample(dev(xyz), dev(feats), dev(labels.astype(np.int32)), 0.15, 3.0)
    same_outputs(got, cp.voxel_downsample(xyz, feats, labels, 0.15, 3.0))


def test_average_downsample(G):
    _, xyz, feats, _ = block(4, 5000, nfeat=9)
    got = G.average_downsample(dev(xyz), dev(feats), 0.1, 1.5)
    same_outputs(got, cp.average_downsample(xyz, feats, 0.1, 1.5))


def test_context_points_pooling(G):
    rng, xyz, feats, _ = block(5, 3000)
    ctx = rng.integers(0, 3000, 700).astype(np.int32)
    got = G.context_points_pooling(dev(xyz), dev(feats), dev(ctx), 0.2, 3.0)
    want = cp.context_points_pooling(xyz, feats, ctx, 0.2, 3.0)
    same_outputs(got, want)
    # the context indices still name the same physical points after the permutation
    assert np.array_equal(want[0][want[-1]], xyz[ctx])


def test_class_pooling(G):
    rng, xyz, feats, labels = block(6, 5000)
    classes = rng.integers(0, 13, 5000).astype(np.int64)
    got = G.class_pooling(dev(xyz), dev(feats), dev(classes), dev(labels), 0.3, 3.0)
    want = cp.class_pooling(xyz, feats, classes, labels, 0.3, 3.0)
    same_outputs(got, want)
    # no voxel mixes two classes (the point of the op, graph_pooling_layer.py:230-257)
    pcls = classes[cp._Level(xyz, 0.3, 3.0, classes.astype(np.int32)).o2p]
    vc = want[-1]
    assert all(len(set(pcls[vc == v])) == 1 for v in range(0, len(want[-3]), 37))


def check_hierarchy(xs, ds, vl, vb, vc):
    """check_vidxs / check_dxyzs at every level + the repermutation kept every coarse point's run of
    fine points in the coarse level's (new) order: centroid(level k) == points of level k+1, bit for bit."""
    for k in range(len(ds)):
        cp.check_vidxs(vl[k], vb[k], vc[k], xs[k].shape[0])
        assert len(vl[k]) == xs[k + 1].shape[0]
        cp.check_dxyzs(xs[k], xs[k + 1], ds[k], vc[k])
        assert np.array_equal(cp.centroid(xs[k], vl[k], vb[k]), xs[k + 1])


@pytest.mark.parametrize("pn", [8000, 700])
def test_points_pooling_two_layers(G, pn):
    _, xyz, feats, labels = block(pn + 1, pn)
    got = G.points_pooling_two_layers(dev(xyz), dev(feats), dev(labels), 0.15, 0.45, 3.0)
    want = cp.points_pooling_two_layers(xyz, feats, labels, 0.15, 0.45, 3.0)
    same_outputs(got, want)
    xs, ds, _, _, vl, vb, vc = want
    check_hierarchy(xs, ds, vl, vb, vc)


def test_context_points_pooling_two_layers(G):
    rng, xyz, feats, labels = block(8, 6000)
    ctx = rng.integers(0, 999, 6000).astype(np.int32)  # one context id per point, carried as a column
    got = G.context_points_pooling_two_layers(dev(xyz), dev(feats), dev(labels), dev(ctx), 0.15, 0.45, 3.0)
    want = cp.context_points_pooling_two_layers(xyz, feats, labels, ctx, 0.15, 0.45, 3.0)
    same_outputs(got, want)
    # labels / features / ctx ids stay attached to their point through both permutations
    key = {tuple(p): (l, c) for p, l, c in zip(map(tuple, xyz), labels, ctx)}
    for p, l, c in list(zip(map(tuple, want[0][0]), want[3], want[7]))[::97]:
        assert key[p] == (l, c)


@pytest.mark.parametrize("pn,voxels", [(9000, (0.15, 0.45, 0.9)), (1200, (0.1, 0.3, 1.5))])
def test_points_pooling_three_layers(G, pn, voxels):
    """The second repermutation chain (graph_pooling_layer.py:307-319) exists only here."""
    _, xyz, feats, labels = block(pn + 2, pn)
    got = G.points_pooling_three_layers(dev(xyz), dev(feats), dev(labels), *voxels, 3.0)
    want = cp.points_pooling_three_layers(xyz, feats, labels, *voxels, 3.0)
    same_outputs(got, want)
    xs, ds, f, lab, vl, vb, vc = want
    assert len(xs) == 4 and len(ds) == 3
    check_hierarchy(xs, ds, vl, vb, vc)
    # features followed their points
    key = {tuple(p): tuple(r) for p, r in zip(map(tuple, xyz), map(tuple, feats))}
    for p, r in list(zip(map(tuple, xs[0]), map(tuple, f)))[::131]:
        assert key[p] == r


def test_batched_two_layers_equals_per_block_oracle(G):
    """block_ids extension: the block-diagonal batch equals the reference composite run once per block,
    indices offset by the block's first point / voxel."""
    from pointcloudsegmentation_b200.synthetic import s3dis_batch

    d = s3dis_batch(9, 4, 1500, 2500)
    bids = d["block_ids"]
    got = G.points_pooling_two_layers(dev(d["xyzs"]), dev(d["feats"]), dev(d["labels"]), 0.15, 0.45, 3.0,
                                      block_ids=dev(bids))
    xs, ds, f, lab, vl, vb, vc, gb = [[t.cpu().numpy() for t in x] if isinstance(x, (list, tuple)) else x.cpu().numpy()
                                      for x in got]
    off = [0, 0, 0]
    for b in range(4):
        m = bids == b
        w = cp.points_pooling_two_layers(d["xyzs"][m], d["feats"][m], d["labels"][m], 0.15, 0.45, 3.0)
        n = [w[0][k].shape[0] for k in range(3)]
        for k in range(3):
            assert np.array_equal(xs[k][off[k]:off[k] + n[k]], w[0][k])
            assert np.all(gb[k][off[k]:off[k] + n[k]] == b)
        for k in range(2):
            assert np.array_equal(ds[k][off[k]:off[k] + n[k]], w[1][k])
            assert np.array_equal(vl[k][off[k + 1]:off[k + 1] + n[k + 1]], w[4][k])
            assert np.array_equal(vb[k][off[k + 1]:off[k + 1] + n[k + 1]] - off[k], w[5][k])
            assert np.array_equal(vc[k][off[k]:off[k] + n[k]] - off[k + 1], w[6][k])
        assert np.array_equal(f[off[0]:off[0] + n[0]], w[2]) and np.array_equal(lab[off[0]:off[0] + n[0]], w[3])
        off = [o + k for o, k in zip(off, n)]
    assert off[0] == len(bids)

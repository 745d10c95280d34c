"""GPU: the libPointUtil replacements (block preparation and label interpolation) through the
C ABI vs the CPU restatement (oracle/point_util_oracle.py).  libPointUtil is un-vendored, so the reference pins
nothing here ("parity unpinned by the reference"); the bar is the repo's own statement of the semantics
(include/pcseg_ops.h): index outputs bit-exact, float outputs 1e-5 relative (covariances 1e-4 absolute: unit-norm
entries from an fp32 two-pass sum vs float64)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import room_block
from oracle import point_util_oracle as O

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def room(seed, pn):
    rng = np.random.default_rng(seed)
    return room_block(rng, pn).astype(np.float32), rng


@pytest.mark.parametrize("pn,stride", [(1, 0.1), (5000, 0.05), (60000, 0.1), (20000, 3.5)])
def test_grid_downsample(pn, stride):
    from pointcloudsegmentation_b200 import point_util as PU

    x, _ = room(pn, pn)
    got = PU.gridDownsampleGPU(dev(x), stride).cpu().numpy()
    want = O.grid_downsample(x, stride)
    assert np.array_equal(got, want)
    # one point per occupied voxel, each the lowest index of its voxel
    v = np.floor((x - x.min(0)) / np.float32(stride)).astype(np.int64)
    assert len({tuple(r) for r in v}) == len(got) == len({tuple(r) for r in v[got]})


@pytest.mark.parametrize("sn,qn,r", [(4000, 1000, 0.1), (4000, 7, 0.3), (33, 100, 10.0), (5000, 2500, 0.02)])
def test_cross_radius_search(sn, qn, r):
    from pointcloudsegmentation_b200 import point_util as PU

    s, rng = room(sn + qn, sn)
    q = (s[rng.integers(0, sn, qn)] + rng.normal(0, 0.02, (qn, 3))).astype(np.float32)
    got = PU.findNeighborInAnotherCPU(dev(s), dev(q), float(r))
    want = O.cross_search(s, q, r * r)
    for g, w in zip(got, want):
        assert np.array_equal(g.cpu().numpy(), w)


def test_radius_search_from_a_subset_of_the_cloud():
    from pointcloudsegmentation_b200 import point_util as PU

    x, rng = room(3, 6000)
    qi = np.sort(rng.choice(6000, 1500, replace=False)).astype(np.int32)
    got = PU.findNeighborRadiusCPU(dev(x), dev(qi), 0.1)
    want = O.cross_search(x, x[qi], 0.1 * 0.1)
    for g, w in zip(got, want):
        assert np.array_equal(g.cpu().numpy(), w)
    # every query finds itself
    idxs, lens, begs, _ = (t.cpu().numpy() for t in got)
    assert all(qi[i] in idxs[begs[i]:begs[i] + lens[i]] for i in range(0, 1500, 97))


@pytest.mark.parametrize("sn,qn,K", [(3000, 20000, 6), (3000, 1001, 8), (40, 100, 32), (6, 50, 6), (5000, 3, 1)])
def test_knn_exact_and_ordered(sn, qn, K):
    from pointcloudsegmentation_b200 import point_util as PU

    s, rng = room(sn * 7 + K, sn)
    q, _ = room(qn + 1, qn)
    if sn >= 100:  # duplicated sources: equal distances must come out in ascending index
        s[50:60] = s[40:50]
    gi, gd = PU.findNeighborInAnotherCPU(dev(s), dev(q), K, return_dists=True)
    wi, wd = O.knn(s, q, K)
    assert np.array_equal(gi.cpu().numpy(), wi)
    assert np.array_equal(gd.cpu().numpy(), wd)  # same fp32 expression, same contraction: bit-exact


def test_knn_far_queries_fall_through_to_the_unbounded_round():
    from pointcloudsegmentation_b200 import point_util as PU

    s, _ = room(5, 2000)
    q = (s[:300] + np.float32(500.0)).astype(np.float32)  # hundreds of metres from every source
    gi = PU.findNeighborInAnotherCPU(dev(s), dev(q), 6).cpu().numpy()
    assert np.array_equal(gi, O.knn(s, q, 6)[0])


@pytest.mark.parametrize("classes", [8, 13, 21, 40])
def test_interpolate_probs(classes):
    from pointcloudsegmentation_b200 import point_util as PU

    s, rng = room(11, 3000)
    q, _ = room(12, 8000)
    pr = rng.uniform(0, 1, (3000, classes)).astype(np.float32)
    idxs, _ = O.knn(s, q, 6)
    lens = np.full(8000, 6, np.int32)
    lens[::17] = 0  # ragged: some queries have no neighbour
    begs = (np.arange(8000) * 6).astype(np.int32)
    ratio = 1.0 / (2 * 0.075 * 0.075)
    got = PU.interpolateProbsGPU(dev(s), dev(q), dev(pr), dev(idxs.reshape(-1)), dev(lens), dev(begs), ratio).cpu().numpy()
    want = O.interpolate_probs(s, q, pr, idxs.reshape(-1), lens, begs, ratio)
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    assert not got[::17].any()


def test_interpolate_composite_and_exact_hits():
    from pointcloudsegmentation_b200 import point_util as PU

    s, rng = room(21, 4000)
    pr = rng.uniform(0, 1, (4000, 13)).astype(np.float32)
    q = np.concatenate([s[:100], room(22, 5000)[0]])
    got = PU.interpolate(dev(s), dev(pr), dev(q)).cpu().numpy()
    idxs, _ = O.knn(s, q, 6)
    n = q.shape[0]
    want = O.interpolate_probs(s, q, pr, idxs.reshape(-1), np.full(n, 6, np.int32), (np.arange(n) * 6).astype(np.int32),
                               1.0 / (2 * 0.075 * 0.075))
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    assert np.array_equal(got.argmax(1), want.argmax(1)) or (got.argmax(1) != want.argmax(1)).mean() < 1e-3


def test_compute_covars():
    from pointcloudsegmentation_b200 import point_util as PU

    x, rng = room(31, 20000)
    qi = np.sort(rng.choice(20000, 4000, replace=False)).astype(np.int32)
    idxs, lens, begs, _ = O.cross_search(x, x[qi], 0.1 * 0.1)
    got = PU.computeCovarsGPU(dev(x), dev(idxs), dev(lens), dev(begs)).cpu().numpy()
    want = O.compute_covars(x, idxs, lens, begs)
    assert np.abs(got - want).max() <= 1e-4
    assert np.allclose((got * got).sum(1)[lens > 3], 1.0, atol=1e-4)  # unit Frobenius norm
    assert np.allclose(got[:, 1], got[:, 3]) and np.allclose(got[:, 5], got[:, 7])  # symmetric
    # a single-point neighbourhood and an empty one give zeros
    l2 = lens.copy(); l2[0] = 1; l2[1] = 0
    g2 = PU.computeCovarsGPU(dev(x), dev(idxs), dev(l2), dev(begs)).cpu().numpy()
    assert not g2[:2].any()


def test_gather_block_points_and_uniform_sample_block():
    from pointcloudsegmentation_b200 import point_util as PU

    rng = np.random.default_rng(41)
    xyz = (rng.uniform(0, 1, (30011, 3)) * np.array([10.0, 7.0, 3.0])).astype(np.float32)
    mx = xyz.max(0)
    begs = [(x, y) for x in O.block_origins(mx[0], 3.0, 1.5) for y in O.block_origins(mx[1], 3.0, 1.5)]
    idxs, lens, bgs = PU.gatherBlockPointsGPU(dev(xyz), begs, 3.0)
    want = O.gather_block_points(xyz, begs, 3.0)
    idxs, lens, bgs = idxs.cpu().numpy(), lens.cpu().numpy(), bgs.cpu().numpy()
    assert len(want) == len(lens) == len(begs)
    for b, w in enumerate(want):
        assert np.array_equal(idxs[bgs[b]:bgs[b] + lens[b]], w)
    blocks = PU.uniform_sample_block(dev(xyz), 3.0, 1.5, min_pn=3500)
    kept = [w for w in want if len(w) >= 3500]
    assert len(blocks) == len(kept) > 0
    for g, w in zip(blocks, kept):
        assert np.array_equal(g.cpu().numpy(), w)


def test_sample_block_pipeline_matches_the_restatement():
    """s3dis_util.sample_block end to end: two grid down-samples, subset radius search, covariances, block crop."""
    from pointcloudsegmentation_b200 import point_util as PU

    rng = np.random.default_rng(51)
    pts = np.concatenate([room_block(rng, 60000, block=6.0).astype(np.float32) + np.float32(3.0),
                          rng.uniform(0, 255, (60000, 3)).astype(np.float32)], axis=1)
    lbl = rng.integers(0, 13, 60000).astype(np.int64)
    got = PU.sample_block(dev(pts), dev(lbl), 0.1, 3.0, 1.5, 256, covar_ds_stride=0.05, covar_nn_size=0.15)
    want = O.sample_block(pts, lbl, 0.1, 3.0, 1.5, 256, covar_ds_stride=0.05, covar_nn_size=0.15)
    assert len(got[0]) == len(want[0]) > 1
    for k in range(len(want[0])):
        assert np.array_equal(got[0][k].cpu().numpy(), want[0][k])          # xyzs
        assert np.array_equal(got[1][k].cpu().numpy(), want[1][k])          # rgbs
        assert np.abs(got[2][k].cpu().numpy() - want[2][k]).max() <= 1e-4   # covars
        assert np.array_equal(got[3][k].cpu().numpy(), want[3][k])          # labels


def test_sort_voxel():
    from pointcloudsegmentation_b200 import point_util as PU

    x, _ = room(61, 30000)
    sidxs, vlens = PU.sortVoxelGPU(dev(x), 0.2)
    sidxs, vlens = sidxs.cpu().numpy(), vlens.cpu().numpy()
    v = np.floor((x - x.min(0)) / np.float32(0.2)).astype(np.int64)
    key = (v[:, 0] << 32) + (v[:, 1] << 16) + v[:, 2]
    want = np.argsort(key, kind="stable")
    assert np.array_equal(sidxs, want.astype(np.int32))
    assert np.array_equal(vlens, np.unique(key, return_counts=True)[1])

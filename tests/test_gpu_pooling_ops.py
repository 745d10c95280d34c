"""GPU parity, rows a12-a20 (+ a21 composites) of SURVEY.md §8: PoolingIndex ops against the CPU oracle,
bit-exact (integer outputs, single-rounding fp32), plus the structural invariants the reference's own
tests check (PoolingIndex/test_classes_permutation.py:47-75, test_neighborhood.py:34-39,
test_neighborhood_fixed.py:32-47, model_pooling.py:774-799)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import room_block
from oracle import pcs_oracle as orc

pytestmark = pytest.mark.gpu


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def exact(got, want):
    got = got.detach().cpu().numpy()
    assert got.shape == want.shape and got.dtype == want.dtype, (got.shape, want.shape, got.dtype, want.dtype)
    assert np.array_equal(got, want)


@pytest.fixture(scope="module")
def P():
    from pointcloudsegmentation_b200 import pooling_ops
    return pooling_ops


@pytest.mark.parametrize("pn,voxel,block", [(30000, 0.3, 2.0), (4096, 0.15, 3.0), (1000, 0.45, 3.0), (1, 0.2, 3.0)])
def test_voxel_index(P, pn, voxel, block):
    rng = np.random.default_rng(pn)
    pts = rng.uniform(-block / 2, block / 2, (pn, 3)).astype(np.float32)
    pts[:, 2] += block / 2
    # adversarial: coordinates sitting exactly on voxel boundaries of the fp32 grid
    k = min(pn, 64)
    mn = np.float32(-np.float32(block) / np.float32(2.0) - np.float32(1e-3))
    pts[:k, 0] = mn + np.arange(k, dtype=np.float32) * np.float32(voxel)
    exact(P.compute_voxel_index(dev(pts), voxel_len=voxel, block_size=block), orc.voxel_index(pts, voxel, block))


@pytest.mark.parametrize("pn,voxel", [(30000, 0.3), (4096, 0.15), (513, 0.05), (1, 0.3)])
def test_permutation_info(P, pn, voxel):
    rng = np.random.default_rng(pn + 1)
    pts = room_block(rng, pn)
    vidx = orc.voxel_index(pts, voxel, 3.0)
    got = P.compute_permutation_info(dev(vidx))
    want = orc.permutation_info(vidx)
    for g, w in zip(got, want):
        exact(g, w)
    o2p, p2o, lens, begs, cens = [t.cpu().numpy() for t in got]
    # structural checks of PoolingIndex/test_classes_permutation.py:47-75
    assert np.array_equal(np.sort(o2p), np.arange(pn)) and np.array_equal(o2p[p2o], np.arange(pn))
    assert np.array_equal(begs, np.cumsum(lens) - lens) and lens.sum() == pn
    assert np.array_equal(cens, np.repeat(np.arange(len(lens)), lens))


def test_permutation_info_with_class_and_blocks(P):
    rng = np.random.default_rng(5)
    pn = 20000
    pts = room_block(rng, pn)
    vidx = orc.voxel_index(pts, 0.3, 3.0)
    classes = rng.integers(0, 13, pn).astype(np.int32)
    for g, w in zip(P.compute_permutation_info_with_class(dev(vidx), dev(classes)), orc.permutation_info(vidx, classes)):
        exact(g, w)
    bids = rng.integers(0, 7, pn).astype(np.int32)  # any order
    for g, w in zip(P.compute_permutation_info(dev(vidx), block_ids=dev(bids)), orc.permutation_info(vidx, None, bids)):
        exact(g, w)


def test_permutation_info_wide_keys(P):
    """Voxel coordinates beyond 16 bits overflow into the neighbouring key field exactly as the
    reference's shift-add does; the 64-bit sort path must reproduce thrust's order."""
    rng = np.random.default_rng(6)
    vidx = rng.integers(0, 70000, (5000, 3)).astype(np.int32)
    vidx[:50, 0] = rng.integers(1 << 20, 1 << 30, 50)
    for g, w in zip(P.compute_permutation_info(dev(vidx)), orc.permutation_info(vidx)):
        exact(g, w)


@pytest.mark.parametrize("pn,ps,dt", [(50, 11, np.float32), (4096, 3, np.float32), (3000, 1, np.int32), (1025, 130, np.float32)])
def test_permutate_feature(P, pn, ps, dt):
    rng = np.random.default_rng(pn)
    o2p = rng.permutation(pn).astype(np.int32)
    p2o = np.empty_like(o2p); p2o[o2p] = np.arange(pn, dtype=np.int32)
    feats = (rng.uniform(-1, 1, (pn, ps)) * 100).astype(dt)
    x = dev(feats)
    if dt == np.float32:
        x.requires_grad_(True)
    y = P.permutate_feature(x, dev(o2p), dev(p2o))
    exact(y, orc.row_gather(feats, o2p))
    if dt == np.float32:
        d = rng.uniform(-1, 1, (pn, ps)).astype(np.float32)
        y.backward(dev(d))
        exact(x.grad, orc.row_gather(d, p2o))  # PermutateFeatureBackward(d, back_idxs)


def test_diff_xyz_has_no_gradient(P):
    rng = np.random.default_rng(3)
    lens = rng.integers(1, 9, 500).astype(np.int32)
    cens = orc.cidxs_from_lens(lens)
    xyz = rng.uniform(-1, 1, (len(cens), 3)).astype(np.float32)
    c = rng.uniform(-1, 1, (500, 3)).astype(np.float32)
    x = dev(xyz).requires_grad_(True)
    y = P.compute_diff_xyz(x, dev(c), dev(cens))
    exact(y, orc.diff_xyz(xyz, c, cens))
    assert not y.requires_grad  # graph_pooling_layer.py:16-18


@pytest.mark.parametrize("pn", [30000, 2048])
def test_repermutation_info(P, pn):
    rng = np.random.default_rng(pn)
    pts = room_block(rng, pn)
    o2p1, p2o1, lens1, begs1, cens1 = orc.permutation_info(orc.voxel_index(pts, 0.15, 3.0))
    pn2 = len(lens1)
    o2p2 = rng.permutation(pn2).astype(np.int32)
    got = P.compute_repermutation_info(dev(o2p2), dev(lens1), dev(begs1), dev(cens1))
    want = orc.repermutation_info(o2p2, lens1, begs1, pn)
    for g, w in zip(got, want):
        exact(g, w)


def inside(pts, idxs, cens, rmin, rmax):
    d = np.linalg.norm(pts[idxs].astype(np.float64) - pts[cens].astype(np.float64), axis=1)
    return np.all(d < rmax + 1e-6) and np.all(d > rmin - 1e-6)


@pytest.mark.parametrize("pn,r", [(1024, 0.15), (4096, 0.1), (5000, 0.3), (33, 0.5), (1, 0.1), (2100, 0.05)])
def test_search_neighborhood(pn, r):
    from pointcloudsegmentation_b200.graph_layers import search_neighborhood, search_neighborhood_range

    rng = np.random.default_rng(pn)
    pts = room_block(rng, pn) if pn > 100 else rng.uniform(0, 1, (pn, 3)).astype(np.float32)
    got = search_neighborhood(dev(pts), r)
    want = orc.search(pts, r * r)
    for g, w in zip(got, want):
        exact(g, w)
    idxs, lens, begs, cens = want
    assert inside(pts, idxs, cens, 0.0, r)
    got = search_neighborhood_range(dev(pts), r, r * 1.5)
    want = orc.search(pts, (r * 1.5) * (r * 1.5), r * r)
    for g, w in zip(got, want):
        exact(g, w)
    assert inside(pts, want[0], want[3], r, r * 1.5)  # PoolingIndex/test_neighborhood.py:34-39


def test_search_boundary_pairs_bit_exact():
    """Lattice points at distances that sit exactly on / one ulp around r: the predicate must be the
    reference's fp32 FMA chain (SearchNeighborhood.cu:24 as nvcc contracts it)."""
    from pointcloudsegmentation_b200.graph_layers import search_neighborhood

    g = np.arange(12, dtype=np.float32) * np.float32(0.05)
    pts = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3).astype(np.float32)
    pts = pts[np.random.default_rng(0).permutation(len(pts))]
    for r in (0.05, 0.1, 0.15, np.sqrt(0.0075)):
        got = search_neighborhood(dev(pts), float(r))
        want = orc.search(pts, float(r) * float(r))
        for a, b in zip(got, want):
            exact(a, b)


@pytest.mark.parametrize("pn,K", [(4096, 10), (500, 3), (1500, 40)])
def test_search_fixed(pn, K):
    from pointcloudsegmentation_b200.graph_layers import search_neighborhood_fixed, search_neighborhood_fixed_range

    rng = np.random.default_rng(pn + K)
    pts = room_block(rng, pn)
    for g, w in zip(search_neighborhood_fixed(dev(pts), 0.1, K), orc.search_fixed(pts, 0.1 * 0.1, K)):
        exact(g, w)
    for g, w in zip(search_neighborhood_fixed_range(dev(pts), 0.1, 0.2, K),
                    orc.search_fixed(pts, 0.2 * 0.2, K, 0.1 * 0.1)):
        exact(g, w)


def test_search_batched_blocks_equals_per_block():
    """block_begs extension == running the reference op once per block and offsetting the indices."""
    from pointcloudsegmentation_b200.graph_layers import search_neighborhood

    rng = np.random.default_rng(11)
    sizes = [1500, 1, 2200, 700]
    blocks = [room_block(rng, n) if n > 1 else np.zeros((1, 3), np.float32) for n in sizes]
    pts = np.concatenate(blocks)
    bb = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    got = search_neighborhood(dev(pts), 0.12, block_begs=dev(bb))
    want = orc.search(pts, 0.12 * 0.12, None, bb)
    for g, w in zip(got, want):
        exact(g, w)
    per = [orc.search(b, 0.12 * 0.12) for b in blocks]
    assert np.array_equal(want[0], np.concatenate([p[0] + o for p, o in zip(per, bb[:-1])]))
    assert np.array_equal(want[1], np.concatenate([p[1] for p in per]))


@pytest.mark.gpu
@pytest.mark.parametrize("pad_k", [None, 0, 1, 7, 4096])
@pytest.mark.parametrize("batched", [False, True])
def test_search_multi_equals_one_search_per_graph(batched, pad_k):
    """csrc/search_multi.cu: every graph of a one-traversal multi search is bit-identical to the single search of a18/a19
    (which the libref parity tests pin to the reference), for 1..5 graphs (5 = a split call), batched or not; with the
    padded hit lists of the count pass (pad_k: default, off, shorter than most rows, shorter than some, longer than all)."""
    import torch

    from pointcloudsegmentation_b200 import ops as P
    from pointcloudsegmentation_b200.synthetic import s3dis_batch

    d = s3dis_batch(17, 3, 1800, 2600)
    x = torch.from_numpy(d["xyzs"]).cuda()
    if batched:
        bids = torch.from_numpy(d["block_ids"]).cuda()
        o2p = P.compute_permutation_info(P.compute_voxel_index(x, 0.15, 3.0), block_ids=bids)[0]
        x = P.permutate_feature(x, o2p)
        counts = torch.bincount(bids.long(), minlength=3)
        bb = torch.cat([torch.zeros(1, dtype=torch.int64, device="cuda"), counts.cumsum(0)]).to(torch.int32)
        kw = dict(block_begs=bb, max_block_pn=int(counts.max()))
    else:
        x = x[: 2000].contiguous()
        kw = {}
    specs = [(None, 0.15 ** 2), (0.15 ** 2, 0.2 ** 2), (0.1 ** 2, 0.15 ** 2), (None, 0.1 ** 2), (0.3 ** 2, 0.45 ** 2)]
    for n in (1, 2, 3, 4, 5):
        got = P.search_neighborhood_multi(x, specs[:n], pad_k=pad_k, **kw)
        assert len(got) == n
        for (lo, hi), g in zip(specs[:n], got):
            want = (P.search_neighborhood_brute_force(x, hi, **kw) if lo is None else
                    P.search_neighborhood_brute_force_range(x, lo, hi, **kw))
            for a, b in zip(g, want):
                assert torch.equal(a, b)
    assert P.search_neighborhood_multi(x[:0].contiguous(), specs[:2])[1][0].numel() == 0

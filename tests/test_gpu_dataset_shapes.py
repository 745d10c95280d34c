"""GPU: the op path at the other BASELINE.json dataset shapes — ModelNet40 (32 clouds x 1024 points, batched
block-diagonally) against the CPU oracle cloud by cloud, and a Semantic3D-sized dense block (~1M points,
too large for the oracle) through size-independent properties: permutation o inverse = identity, runs
contiguous, every listed neighbour inside the radius, the graph symmetric, a sampled subset complete against
brute force, scatter -> gather adjointness."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import pcs_oracle as orc
from pointcloudsegmentation_b200 import ops as P
from pointcloudsegmentation_b200.graph_layers import points_pooling_two_layers, search_neighborhood
from pointcloudsegmentation_b200.synthetic import modelnet_batch, semantic3d_block

pytestmark = pytest.mark.gpu


def test_modelnet40_batch_matches_oracle_per_cloud():
    B, pn = 32, 1024
    d = modelnet_batch(0, B, pn)
    x = torch.from_numpy(d["xyzs"]).cuda()
    bb = torch.arange(B + 1, dtype=torch.int32, device="cuda") * pn
    for r in (0.05, 0.1, 0.3):
        nidxs, nlens, nbegs, ncens = search_neighborhood(x, r, block_begs=bb, max_block_pn=pn)
        hn, hl, hb = nidxs.cpu().numpy(), nlens.cpu().numpy(), nbegs.cpu().numpy()
        for c in (0, 13, 31):  # oracle cloud by cloud (indices local to the cloud)
            wi, wl, wb, _ = orc.search(d["xyzs"][c * pn:(c + 1) * pn], np.float32(r) * np.float32(r))
            sl = slice(c * pn, (c + 1) * pn)
            assert np.array_equal(hl[sl], wl)
            e0 = hb[c * pn]
            assert np.array_equal(hb[sl] - e0, wb)
            assert np.array_equal(hn[e0:e0 + wi.shape[0]] - c * pn, wi)
    # features through scatter -> max-gather on the r = 0.1 graph of the whole batch vs the oracle
    nidxs, nlens, nbegs, ncens = search_neighborhood(x, 0.1, block_begs=bb, max_block_pn=pn)
    f = torch.rand(B * pn, 64, device="cuda") * 2 - 1
    s = P.neighbor_scatter(f, nidxs, nlens, nbegs, True)
    g, mi = P.neighbor_max_feat_gather(s, nlens, nbegs)
    ws = orc.neighbor_scatter(f.cpu().numpy(), nidxs.cpu().numpy(), nlens.cpu().numpy(), nbegs.cpu().numpy(), True)
    wg, wmi = orc.max_feat_gather(ws, nlens.cpu().numpy(), nbegs.cpu().numpy())
    assert np.array_equal(s.cpu().numpy(), ws)
    assert np.array_equal(g.cpu().numpy(), wg) and np.array_equal(mi.cpu().numpy(), wmi)


@pytest.fixture(scope="module")
def sem3d():
    pts = semantic3d_block(0, 1_000_000)
    x = torch.from_numpy(pts).cuda()
    f = torch.rand(x.shape[0], 4, device="cuda")
    lab = torch.zeros(x.shape[0], dtype=torch.int64, device="cuda")
    out = points_pooling_two_layers(x, f, lab, 0.45, 1.5, 80.0)
    return x, out


def test_semantic3d_block_pooling_properties(sem3d):
    x, (xs, dxs, f, lab, vl, vb, vc) = sem3d
    pn = x.shape[0]
    assert xs[0].shape[0] == pn and xs[1].shape[0] == vl[0].shape[0] and xs[2].shape[0] == vl[1].shape[0]
    for lens, begs, cens, n in ((vl[0], vb[0], vc[0], pn), (vl[1], vb[1], vc[1], xs[1].shape[0])):
        assert int(lens.sum()) == n and int(lens.min()) >= 1
        assert torch.equal(begs, torch.cumsum(lens, 0, dtype=torch.int32) - lens)
        assert torch.equal(cens, torch.repeat_interleave(torch.arange(lens.shape[0], device="cuda", dtype=torch.int32),
                                                         lens.long()))
    # the level-0 points are a permutation of the input: same multiset of coordinates
    a = torch.sort(x[:, 0] * 7 + x[:, 1] * 13 + x[:, 2] * 29).values
    b = torch.sort(xs[0][:, 0] * 7 + xs[0][:, 1] * 13 + xs[0][:, 2] * 29).values
    assert torch.equal(a, b)
    # every point lies within its voxel's extent of the voxel centroid
    d = (xs[0] - xs[1][vc[0].long()]).abs().max()
    assert float(d) <= 0.45 + 1e-5


@pytest.mark.parametrize("r", [0.3, 0.2])
def test_semantic3d_block_search_properties(sem3d, r):
    _, (xs, *_rest) = sem3d
    x = xs[0]
    pn = x.shape[0]
    nidxs, nlens, nbegs, ncens = search_neighborhood(x, r)
    E = nidxs.shape[0]
    assert int(nlens.sum()) == E and torch.equal(nbegs, torch.cumsum(nlens, 0, dtype=torch.int32) - nlens)
    r2 = float(np.float32(r) * np.float32(r))
    d2 = ((x[nidxs.long()] - x[ncens.long()]) ** 2).sum(1)
    assert float(d2.max()) < r2 * (1 + 1e-5)            # every listed neighbour inside the radius
    assert int(nlens.min()) >= 1                         # self edge always present
    # ascending neighbour index inside each row
    same_row = ncens[1:] == ncens[:-1]
    assert bool((nidxs[1:][same_row] > nidxs[:-1][same_row]).all())
    # symmetric: the multiset of (i, j) equals the multiset of (j, i)
    k1 = ncens.long() * pn + nidxs.long()
    k2 = nidxs.long() * pn + ncens.long()
    assert torch.equal(torch.sort(k1).values, torch.sort(k2).values)
    # complete on a sample of centres against brute force (pairs within 1e-6 of the boundary excluded)
    g = torch.Generator(device="cuda").manual_seed(1)
    sample = torch.randint(0, pn, (512,), device="cuda", generator=g)
    dd = torch.cdist(x[sample].double(), x.double()) ** 2
    inside = (dd < r2 * (1 - 1e-6)).sum(1)
    maybe = (dd < r2 * (1 + 1e-6)).sum(1)
    got = nlens[sample].long()
    assert bool((got >= inside).all()) and bool((got <= maybe).all())


def test_semantic3d_block_scatter_gather_adjoint(sem3d):
    _, (xs, *_rest) = sem3d
    x = xs[0]
    nidxs, nlens, nbegs, ncens = search_neighborhood(x, 0.2)
    g = torch.Generator(device="cuda").manual_seed(2)
    f = torch.rand(x.shape[0], 16, device="cuda", generator=g)
    w = torch.rand(nidxs.shape[0], 16, device="cuda", generator=g)
    for diff in (False, True):
        lhs = (P.neighbor_scatter(f, nidxs, nlens, nbegs, diff).double() * w.double()).sum()
        rhs = (f.double() * P.neighbor_gather(w, nidxs, nlens, nbegs, diff).double()).sum()
        assert abs(float(lhs - rhs)) <= 1e-6 * abs(float(lhs))


def test_semantic3d_block_preparation_properties():
    """Block preparation of a 1M-point Semantic3D tile (what bench.py --dataset semantic3d runs per step) at full size,
    through size-independent properties — the CPU restatement's searches are O(n^2): one point per occupied voxel,
    each the lowest index of its voxel; the centre-vs-cloud radius graph complete and sound on a sample against brute
    force; k-NN rows sorted, exact on a sample; interpolation of a constant field is the constant."""
    from pointcloudsegmentation_b200 import point_util as PU

    x = torch.from_numpy(semantic3d_block(3, 1_000_000)).cuda()
    x = x - x.min(dim=0).values
    pn = x.shape[0]
    # voxel sort + grid down-sample
    sidxs, vlens = PU.sortVoxelGPU(x, 0.15)
    assert int(vlens.sum()) == pn and torch.equal(torch.sort(sidxs).values, torch.arange(pn, device="cuda", dtype=torch.int32))
    idxs = PU.gridDownsampleGPU(x, 0.15)
    assert idxs.shape[0] == vlens.shape[0]
    d = x - x.min(dim=0).values
    v = torch.floor(d / torch.full_like(d, 0.15)).long()  # tensor / tensor: an IEEE divide (a scalar divisor is folded into a reciprocal multiply)
    key = (v[:, 0] << 32) + (v[:, 1] << 16) + v[:, 2]
    assert torch.unique(key).shape[0] == idxs.shape[0] and bool((key[idxs.long()][1:] > key[idxs.long()][:-1]).all())
    first = torch.full((int(key.max()) + 1,), pn, device="cuda", dtype=torch.int64) if int(key.max()) < 1 << 28 else None
    if first is not None:
        first.scatter_reduce_(0, key, torch.arange(pn, device="cuda"), reduce="amin")
        assert torch.equal(first[key[idxs.long()]], idxs.long())
    # centre-vs-cloud radius graph on the voxel-sorted cloud
    xs = x[sidxs.long()].contiguous()
    cen = PU.gridDownsampleGPU(xs, 0.15)
    nidxs, nlens, nbegs, ncens = PU.findNeighborRadiusCPU(xs, cen, 0.25)
    assert int(nlens.sum()) == nidxs.shape[0] and torch.equal(nbegs, torch.cumsum(nlens, 0, dtype=torch.int32) - nlens)
    r2 = float(np.float32(0.25 * 0.25))
    q = xs[cen.long()]
    d2 = ((xs[nidxs.long()] - q[ncens.long()]) ** 2).sum(1)
    assert float(d2.max()) < r2 * (1 + 1e-5) and int(nlens.min()) >= 1
    g = torch.Generator(device="cuda").manual_seed(4)
    sample = torch.randint(0, cen.shape[0], (256,), device="cuda", generator=g)
    dd = torch.cdist(q[sample].double(), xs.double()) ** 2
    got = nlens[sample].long()
    assert bool((got >= (dd < r2 * (1 - 1e-6)).sum(1)).all()) and bool((got <= (dd < r2 * (1 + 1e-6)).sum(1)).all())
    # k-NN of the full cloud in the centres: sorted rows, exact on the sample
    knn, kd = PU.findNeighborInAnotherCPU(q.contiguous(), xs, 8, return_dists=True)
    assert bool((kd[:, 1:] >= kd[:, :-1]).all()) and int(knn.min()) >= 0 and int(knn.max()) < q.shape[0]
    qs = torch.randint(0, pn, (256,), device="cuda", generator=g)
    ref = torch.topk(torch.cdist(xs[qs].double(), q.double()) ** 2, 8, dim=1, largest=False).values
    assert torch.allclose(kd[qs].double(), ref, rtol=1e-5, atol=1e-9)
    # Gaussian interpolation reproduces a constant field and stays inside the convex hull of the sources' values
    probs = torch.rand(q.shape[0], 8, device="cuda", generator=g)
    const = torch.full((q.shape[0], 8), 0.37, device="cuda")
    lens8 = torch.full((pn,), 8, dtype=torch.int32, device="cuda")
    begs8 = torch.arange(pn, dtype=torch.int32, device="cuda") * 8
    out_c = PU.interpolateProbsGPU(q.contiguous(), xs, const, knn.reshape(-1), lens8, begs8, 1.0 / (2 * 0.15 * 0.15))
    assert torch.allclose(out_c, torch.full_like(out_c, 0.37), atol=1e-6)
    out = PU.interpolateProbsGPU(q.contiguous(), xs, probs, knn.reshape(-1), lens8, begs8, 1.0 / (2 * 0.15 * 0.15))
    nb = probs[knn.long()]  # [pn, 8, classes]
    assert bool((out <= nb.max(1).values + 1e-6).all()) and bool((out >= nb.min(1).values - 1e-6).all())

"""GPU parity, rows a1-a11 of SURVEY.md §8: the sm_100a kernels (through the reference-named Python
ops -> ctypes -> C ABI) against the CPU oracle on seeded inputs and against the golden vectors the
reference's numpy oracles produced.  Shapes follow the reference tests' regimes
(tf_ops/test/test_neighbor_scatter.py:238-259).  Bars: bit-exact for copies / single-rounding results /
indices / argmax; fp32 sums within 1e-5 relative of the float64 oracle (north_star), which also
satisfies the reference's mean<=1e-5, max<=1e-4 absolute bar."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import GOLDEN, random_csr
from oracle import pcs_oracle as orc

pytestmark = pytest.mark.gpu

REGIMES = [(57, 33), (300, 64), (1030, 31), (1024, 320), (401, 1030), (5125, 77), (64, 3), (7, 1), (2000, 128)]


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def sums_close(got, want64, scale=None):
    """|got - want| <= 1e-5 * max(|want|, scale): fp32 sums vs float64 oracle (north_star 1e-5 rel)."""
    got = got.detach().cpu().numpy().astype(np.float64)
    assert got.shape == want64.shape
    if got.size == 0:
        return
    ref = np.maximum(np.abs(want64), 1.0 if scale is None else scale)
    err = np.abs(got - want64) / ref
    assert err.max() <= 1e-5, err.max()
    d = np.abs(got - want64)
    assert d.mean() <= 1e-5 and d.max() <= 1e-4  # the reference tests' own bar


def exact(got, want):
    got = got.detach().cpu().numpy()
    assert got.shape == want.shape and got.dtype == want.dtype, (got.shape, want.shape, got.dtype, want.dtype)
    assert np.array_equal(got, want)


@pytest.fixture(scope="module")
def ops():
    from pointcloudsegmentation_b200 import neighbor_ops
    return neighbor_ops


@pytest.mark.parametrize("pn,fd", REGIMES)
@pytest.mark.parametrize("use_diff", [False, True])
def test_scatter_gather(ops, pn, fd, use_diff):
    rng = np.random.default_rng(pn * 1000 + fd)
    nidxs, lens, bgs, cens = random_csr(rng, pn, 10)
    feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
    ds = rng.uniform(-1, 1, (len(nidxs), fd)).astype(np.float32)
    a = (dev(nidxs), dev(lens), dev(bgs))
    x = dev(feats).requires_grad_(True)
    y = ops.neighbor_scatter(x, *a, use_diff=use_diff)
    exact(y, orc.neighbor_scatter(feats, nidxs, lens, bgs, use_diff))  # copy / one subtract: bit-exact
    y.backward(dev(ds))
    sums_close(x.grad, orc.neighbor_gather(ds, nidxs, lens, bgs, use_diff))
    # determinism: the gather is an ordered segmented reduction, not atomics
    g2 = ops.neighbor_gather(dev(ds), *a, use_diff=use_diff)
    assert torch.equal(g2, x.grad)


@pytest.mark.parametrize("pn,fd", REGIMES)
def test_sum_feat_gather_scatter(ops, pn, fd):
    rng = np.random.default_rng(pn * 7 + fd)
    _, lens, bgs, cens = random_csr(rng, pn, 5)
    feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
    ds = rng.uniform(-1, 1, (len(cens), fd)).astype(np.float32)
    a = (dev(cens), dev(lens), dev(bgs))
    x = dev(feats).requires_grad_(True)
    y = ops.neighbor_sum_feat_scatter(x, *a)
    exact(y, orc.row_gather(feats, cens))
    y.backward(dev(ds))
    sums_close(x.grad, orc.segment_sum(ds, lens, bgs))
    e = dev(ds).requires_grad_(True)
    s = ops.neighbor_sum_feat_gather(e, *a)
    sums_close(s, orc.segment_sum(ds, lens, bgs))
    s.backward(dev(feats))
    exact(e.grad, orc.row_gather(feats, cens))


@pytest.mark.parametrize("pn2,fd,lo,hi", [(30, 19, 3, 5), (500, 64, 0, 10), (1000, 130, 1, 40), (64, 3, 0, 3),
                                           (3000, 32, 1, 70), (100, 1, 0, 5)])
def test_max_feat_gather_scatter(ops, pn2, fd, lo, hi):
    rng = np.random.default_rng(pn2 + fd)
    vlens = rng.integers(lo, hi, pn2).astype(np.int32)
    vbegs = orc.bgs_from_lens(vlens)
    pn1 = int(vlens.sum())
    # quantised values => many exact ties: the FIRST maximum must win (strict '<', TFNeighborKernelNew.cu:111)
    feats = (rng.integers(-4, 5, (pn1, fd)) / 4.0).astype(np.float32)
    dg = rng.uniform(-1, 1, (pn2, fd)).astype(np.float32)
    x = dev(feats).requires_grad_(True)
    g, mi = ops.neighbor_max_feat_gather(x, dev(vlens), dev(vbegs))
    og, omi = orc.max_feat_gather(feats, vlens, vbegs)
    exact(g, og)
    exact(mi, omi)
    g.backward(dev(dg))
    exact(x.grad, orc.max_feat_scatter(dg, omi, vbegs, vlens, pn1))


def test_max_gather_nan_rule(ops):
    """A NaN is selected only when it is the first row of the segment (strict '<' is false for NaN)."""
    f = np.array([[np.nan, 1.0], [2.0, np.nan], [1.0, 0.5]], np.float32)
    g, mi = ops.neighbor_max_feat_gather(dev(f), dev(np.array([3], np.int32)), dev(np.array([0], np.int32)))
    og, omi = orc.max_feat_gather(f, np.array([3], np.int32), np.array([0], np.int32))
    assert np.array_equal(mi.cpu().numpy(), omi)
    assert np.array_equal(g.cpu().numpy(), og, equal_nan=True)


@pytest.mark.parametrize("pn,m", [(57, 26), (1000, 26), (333, 7), (2048, 64), (10, 1)])
def test_location_weight_sum(ops, pn, m):
    rng = np.random.default_rng(pn + m)
    _, lens, bgs, _ = random_csr(rng, pn, 16)
    csum = int(lens.sum())
    lw = rng.uniform(0, 1, (csum, m)).astype(np.float32)
    d = rng.uniform(-1, 1, (pn, m)).astype(np.float32)
    x = dev(lw).requires_grad_(True)
    s = ops.location_weight_sum(x, dev(lens), dev(bgs))
    sums_close(s, orc.segment_sum(lw, lens, bgs))
    s.backward(dev(d))
    exact(x.grad, orc.locw_sum_backward(d, lens, bgs, csum))


@pytest.mark.parametrize("pn,m,ofn", [(57, 26, 8), (300, 26, 32), (129, 5, 3), (64, 26, 64), (100, 3, 1), (40, 26, 16)])
def test_location_weight_feat_sum(ops, pn, m, ofn):
    rng = np.random.default_rng(pn + m + ofn)
    _, lens, bgs, _ = random_csr(rng, pn, 14)
    csum = int(lens.sum())
    lw = rng.uniform(0, 1, (csum, m)).astype(np.float32)
    tf = rng.uniform(-1, 1, (csum, m, ofn)).astype(np.float32)
    d = rng.uniform(-1, 1, (pn, m, ofn)).astype(np.float32)
    xl, xt = dev(lw).requires_grad_(True), dev(tf).requires_grad_(True)
    s = ops.location_weight_feat_sum(xl, xt, dev(lens), dev(bgs))
    sums_close(s, orc.locw_feat_sum(lw, tf, lens, bgs))
    s.backward(dev(d))
    odlw, odt = orc.locw_feat_sum_backward(lw, tf, d, lens, bgs)
    exact(xt.grad, odt)  # one fp32 multiply per element
    sums_close(xl.grad, odlw)


@pytest.mark.parametrize("pn,r2", [(40, 0.09), (700, 0.01), (2000, 0.004)])
def test_eliminate_center(ops, pn, r2):
    rng = np.random.default_rng(pn)
    pts = rng.uniform(0, 1, (pn, 3)).astype(np.float32)
    nidxs, lens, bgs, _ = orc.search(pts, r2)  # every row contains its centre exactly once
    got = ops.eliminate_center(dev(nidxs), dev(lens), dev(bgs))
    want = orc.eliminate_center(nidxs, lens, bgs)
    for g, w in zip(got, want):
        exact(g, w)


# ---- golden vectors (reference numpy oracles, tests/golden/make_golden.py) ------------------------------
def gclose(got, want):
    d = np.abs(got.detach().cpu().numpy().astype(np.float64) - want)
    assert d.mean() <= 1e-5 and d.max() <= 1e-4  # tf_ops/test/test_neighbor_scatter.py:194-199


@pytest.mark.parametrize("tag", ["small", "wide", "odd"])
def test_golden_neighbor_scatter(ops, tag):
    g = np.load(GOLDEN / f"neighbor_scatter_{tag}.npz")
    a = (dev(g["nidxs"]), dev(g["lens"]), dev(g["bgs"]))
    for use_diff, k in [(False, "copy"), (True, "diff")]:
        x = dev(g["feats"]).requires_grad_(True)
        y = ops.neighbor_scatter(x, *a, use_diff=use_diff)
        gclose(y, g[f"{k}_fwd"])
        y.backward(dev(g["dsfeats"]))
        gclose(x.grad, g[f"{k}_bwd"])


@pytest.mark.parametrize("tag", ["small", "wide"])
def test_golden_sum_scatter(ops, tag):
    g = np.load(GOLDEN / f"sum_scatter_{tag}.npz")
    x = dev(g["feats"]).requires_grad_(True)
    y = ops.neighbor_sum_feat_scatter(x, dev(g["cens"]), dev(g["lens"]), dev(g["bgs"]))
    gclose(y, g["fwd"])
    y.backward(dev(g["dsfeats"]))
    gclose(x.grad, g["bwd"])


@pytest.mark.parametrize("tag", ["pool", "ragged"])
def test_golden_max_pool_unpool(tag):
    from pointcloudsegmentation_b200.graph_layers import graph_pool_for_test, graph_unpool

    g = np.load(GOLDEN / f"max_pool_{tag}.npz")
    x = dev(g["ifeats"]).requires_grad_(True)
    p, mi = graph_pool_for_test(x, dev(g["vlens"]), dev(g["vbegs"]))
    gclose(p, g["pfeats"])
    assert np.array_equal(mi.cpu().numpy(), g["max_idxs"])
    p.backward(dev(g["dpfeats"]))
    gclose(x.grad, g["difeats"])
    u = dev(g["up_in"]).requires_grad_(True)
    up = graph_unpool(u, dev(g["vlens"]), dev(g["vbegs"]), dev(g["vcens"]))
    gclose(up, g["up"])
    up.backward(dev(g["dup"]))
    gclose(u.grad, g["dup_in"])


def test_golden_concat_non_center():
    from pointcloudsegmentation_b200.graph_layers import graph_concat_non_center_scatter

    g = np.load(GOLDEN / "concat_non_center.npz")
    x = dev(g["feats"]).requires_grad_(True)
    y = graph_concat_non_center_scatter(x, dev(g["nidxs"]), dev(g["lens"]), dev(g["bgs"]))
    gclose(y, g["fwd"])
    y.backward(dev(g["dfwd"]))
    gclose(x.grad, g["bwd"])


def test_csr_transpose_symmetric_equals_sorted_transpose():
    """Radius graphs are symmetric with ascending rows: the binary-search transpose must equal the radix-sort one
    bit for bit; a non-symmetric CSR must be reported."""
    import numpy as np

    from pointcloudsegmentation_b200 import ops
    from pointcloudsegmentation_b200.graph_layers import search_neighborhood, search_neighborhood_range

    rng = np.random.default_rng(5)
    pts = torch.from_numpy(rng.uniform(-1.5, 1.5, (6000, 3)).astype(np.float32)).cuda()
    for g in (search_neighborhood(pts, 0.2), search_neighborhood_range(pts, 0.1, 0.25)):
        nidxs, nlens, nbegs, ncens = g
        tp, tb = ops.csr_transpose(nidxs.clone(), nlens.shape[0])
        sp, sb = ops.csr_transpose_symmetric(nidxs, nlens, nbegs, ncens)
        assert torch.equal(tp, sp) and torch.equal(tb, sb)
    # not symmetric: drop the mirror of one edge by redirecting it
    nidxs, nlens, nbegs, ncens = search_neighborhood(pts, 0.2)
    bad = nidxs.clone()
    e = int(torch.nonzero(ncens != nidxs)[0])
    bad[e] = ncens[e]  # now a duplicate self edge; the row is no longer a mirror image
    with pytest.raises(ValueError):
        ops.csr_transpose_symmetric(bad, nlens, nbegs, ncens)


@pytest.mark.gpu
@pytest.mark.parametrize("gaps", [False, True])
def test_eliminate_center_irregular_rows(gaps):
    """a11 on rows the radius graphs never produce: no self edge (the last entry is dropped, the reference's
    `pos < len - 1` guard, TFNeighborKernelNew.cu:337-370), several self edges, single-entry rows, a CSR with gaps — the
    edge-parallel tile kernel and the sub-warp ballot kernel both against a direct numpy restatement."""
    import torch

    from pointcloudsegmentation_b200 import ops as P

    rng = np.random.default_rng(5)
    pn = 3000
    lens = rng.integers(1, 40, pn).astype(np.int32)  # rows are never empty: every row loses exactly one slot
    lens[::7] = 1
    stride = lens.astype(np.int64) + (rng.integers(0, 3, pn) if gaps else 0)
    bgs = (np.cumsum(stride) - stride).astype(np.int32)
    csum = int(bgs[-1] + stride[-1])
    nidxs = rng.integers(0, pn, csum).astype(np.int32)
    kind = rng.integers(0, 4, pn)  # 0: no self, 1-2: one self, 3: several
    for i in range(pn):
        row = nidxs[bgs[i]:bgs[i] + lens[i]]
        row[row == i] = (i + 1) % pn
        if kind[i] in (1, 2):
            row[rng.integers(0, lens[i])] = i
        elif kind[i] == 3:
            row[rng.integers(0, lens[i], 3)] = i
    for mode in (1, 0):
        P.set_option("elim_tile", mode)
        try:
            on, ol, ob, oc = P.eliminate_center(torch.from_numpy(nidxs).cuda(), torch.from_numpy(lens).cuda(), torch.from_numpy(bgs).cuda())
        finally:
            P.set_option("elim_tile", 1)
        on, ol, ob, oc = (t.cpu().numpy() for t in (on, ol, ob, oc))
        assert np.array_equal(ol, lens - 1) and np.array_equal(ob, bgs - np.arange(pn, dtype=np.int32))
        for i in range(0, pn, 3):
            row = nidxs[bgs[i]:bgs[i] + lens[i]]
            kept = row[row != i][:max(lens[i] - 1, 0)]
            w = bgs[i] - i
            if len(kept) and w >= 0 and w + len(kept) <= len(on):
                assert np.array_equal(on[w:w + len(kept)], kept), (mode, i)
                assert np.all(oc[w:w + len(kept)] == i)

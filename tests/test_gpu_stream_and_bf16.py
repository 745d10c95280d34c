"""GPU parity for (1) the TMA-staged stream kernels (csrc/seg_stream.cu: segmented sum / max / broadcast for narrow
and odd rows) and (2) the bf16-storage twins of the feature movers (rows a1-a10, a15; VERDICT n1).

fp32: segment sums are SEQUENTIAL fp32 sums in edge order in the reference kernel (TFNeighborKernelNew.cu:19-25),
in oracle.segment_sum_f32 and in both kernel families here, so they are compared BIT-EXACT (not 1e-5); the walker
and the stream kernel are also run on the same input (pcs_set_option seg_stream) and must agree
bit for bit.  Long segments force tiles of several 16 KB chunks; gap CSRs force the per-tile fallback.

bf16: inputs are rounded to bf16 first and the oracle is fed the same (bf16-representable) values.  Copies,
differences (one fp32 subtract, one rounding), products and max / arg-max: bit-exact after rounding the oracle's
fp32 result to bf16 (nearest-even).  Sums: within 1e-2 relative of the float64 oracle (north_star), scale max(|x|,1)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from conftest import random_csr
from oracle import pcs_oracle as orc

pytestmark = pytest.mark.gpu

WIDTHS = [1, 3, 8, 12, 26, 35, 64, 131]


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def bf(a):
    """fp32 array rounded to bf16 (nearest-even), returned as fp32."""
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def to_bf(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).cuda().to(torch.bfloat16)


def same(got, want):
    got = got.detach().float().cpu().numpy() if got.dtype == torch.bfloat16 else got.detach().cpu().numpy()
    assert got.shape == want.shape, (got.shape, want.shape)
    assert np.array_equal(got, want, equal_nan=True), int((got != want).sum())


def close(got, want, tol):
    got = got.detach().float().cpu().numpy().astype(np.float64)
    want = np.asarray(want, np.float64)
    assert got.shape == want.shape
    if got.size:
        assert (np.abs(got - want) / np.maximum(np.abs(want), 1.0)).max() <= tol


@pytest.fixture(scope="module")
def N():
    from pointcloudsegmentation_b200 import ops
    return ops


def long_csr(rng, pn, max_len):
    """segments from 0 to max_len rows (a few very long ones): tiles of many chunks, ragged ends everywhere"""
    lens = rng.integers(0, 12, pn).astype(np.int32)
    lens[rng.integers(0, pn, max(1, pn // 50))] = rng.integers(max_len // 2, max_len, max(1, pn // 50))
    bgs = orc.bgs_from_lens(lens)
    return lens, bgs, orc.cidxs_from_lens(lens)


@pytest.mark.parametrize("C", WIDTHS)
@pytest.mark.parametrize("pn,max_len", [(3000, 40), (700, 3000), (1, 5)])
def test_segment_sum_and_max_bit_exact_both_kernel_families(N, C, pn, max_len):
    rng = np.random.default_rng(C * 1000 + pn)
    lens, bgs, cens = long_csr(rng, pn, max_len)
    E = int(lens.sum())
    x = (rng.integers(-8, 9, (E, C)) / 8.0 + rng.uniform(-1e-3, 1e-3, (E, C))).astype(np.float32)  # near-ties
    dx, dl, db, dc = dev(x), dev(lens), dev(bgs), dev(cens)
    want_sum = orc.segment_sum_f32(x, lens, bgs)
    want_g, want_mi = orc.max_feat_gather(x, lens, bgs)
    outs = []
    for rb in (2, 0):  # stream kernel for every width / walkers only
        N.set_option("seg_stream", rb)
        try:
            s = N.neighbor_sum_feat_gather(dx, dc, dl, db)
            g, mi = N.neighbor_max_feat_gather(dx, dl, db)
        finally:
            N.set_option("seg_stream", 1)
        same(s, want_sum); same(g, want_g); same(mi, want_mi)
        outs.append((s, g, mi))
    for a, b in zip(*outs):
        assert torch.equal(a, b)


@pytest.mark.parametrize("C", WIDTHS)
@pytest.mark.parametrize("pn,max_len", [(3000, 40), (500, 3000)])
def test_broadcast_and_max_scatter_both_kernel_families(N, C, pn, max_len):
    rng = np.random.default_rng(C * 77 + pn)
    lens, bgs, cens = long_csr(rng, pn, max_len)
    E = int(lens.sum())
    f = rng.uniform(-1, 1, (pn, C)).astype(np.float32)
    x = rng.uniform(-1, 1, (E, C)).astype(np.float32)
    _, mi = orc.max_feat_gather(x, lens, bgs)
    want_b = orc.row_gather(f, cens)
    want_m = orc.max_feat_scatter(f, mi, bgs, lens, E)
    df, dl, db, dc, dmi = dev(f), dev(lens), dev(bgs), dev(cens), dev(mi)
    for rb in (2, 0):
        N.set_option("seg_stream", rb)
        try:
            same(N.neighbor_sum_feat_scatter(df, dc, dl, db), want_b)
            same(N.location_weight_sum_backward(torch.empty(E, C, device="cuda"), df, dl, db), want_b)
            same(N.neighbor_max_feat_scatter(df, E, dmi, db, dl), want_m)
        finally:
            N.set_option("seg_stream", 1)


@pytest.mark.parametrize("C", [3, 12, 35])
def test_gap_csr_takes_the_fallback_and_zero_fills(N, C):
    """bgs is NOT the scan of lens (the op signature allows it): sums read the right rows, broadcasts leave the
    reference's memset zeros in the gap rows (TFNeighborKernel.cu:452, TFNeighborKernelNew.cu:204)."""
    rng = np.random.default_rng(C)
    pn = 900
    lens = rng.integers(0, 9, pn).astype(np.int32)
    gaps = (rng.random(pn) < 0.1) * rng.integers(1, 4, pn)
    bgs = (np.cumsum(lens + gaps) - lens).astype(np.int32)  # first segment may start after a gap too
    E = int(bgs[-1] + lens[-1]) + 5  # and rows after the last segment
    x = rng.uniform(-1, 1, (E, C)).astype(np.float32)
    f = rng.uniform(-1, 1, (pn, C)).astype(np.float32)
    dx, df, dl, db = dev(x), dev(f), dev(lens), dev(bgs)
    for rb in (2, 0):
        N.set_option("seg_stream", rb)
        try:
            same(N.neighbor_sum_feat_gather(dx, torch.zeros(E, dtype=torch.int32, device="cuda"), dl, db),
                 orc.segment_sum_f32(x, lens, bgs))
            g, mi = N.neighbor_max_feat_gather(dx, dl, db)
            og, omi = orc.max_feat_gather(x, lens, bgs)
            same(g, og); same(mi, omi)
            same(N.location_weight_sum_backward(torch.empty(E, C, device="cuda"), df, dl, db),
                 orc.locw_sum_backward(f, lens, bgs, E))
            same(N.neighbor_max_feat_scatter(df, E, mi, db, dl), orc.max_feat_scatter(f, omi, bgs, lens, E))
        finally:
            N.set_option("seg_stream", 1)


# ---- bf16 storage ------------------------------------------------------------------------------------
@pytest.mark.parametrize("C", WIDTHS)
def test_bf16_scatter_gather_family(N, C):
    rng = np.random.default_rng(C + 5)
    pn = 1500
    nidxs, lens, bgs, cens = random_csr(rng, pn, 14, 0)
    E = len(nidxs)
    f = bf(rng.uniform(-1, 1, (pn, C)))
    d = bf(rng.uniform(-1, 1, (E, C)))
    bfeat, bd = to_bf(f), to_bf(d)
    dn, dl, db, dc = dev(nidxs), dev(lens), dev(bgs), dev(cens)
    for ud in (False, True):
        out = N.neighbor_scatter(bfeat, dn, dl, db, ud)
        assert out.dtype == torch.bfloat16
        same(out, bf(orc.neighbor_scatter(f, nidxs, lens, bgs, ud)))  # one fp32 subtract, one rounding
        close(N.neighbor_gather(bd, dn, dl, db, ud), orc.neighbor_gather(d, nidxs, lens, bgs, ud), 1e-2)
    close(N.neighbor_sum_feat_gather(bd, dc, dl, db), orc.segment_sum(d, lens, bgs), 1e-2)
    same(N.neighbor_sum_feat_gather(bd, dc, dl, db), bf(orc.segment_sum_f32(d, lens, bgs)))  # fp32 sequential sum, rounded once
    same(N.neighbor_sum_feat_scatter(bfeat, dc, dl, db), orc.row_gather(f, cens))
    g, mi = N.neighbor_max_feat_gather(bd, dl, db)
    og, omi = orc.max_feat_gather(d, lens, bgs)
    same(g, og); same(mi, omi)  # arg-max computed on the bf16 values: indices bit-exact
    same(N.neighbor_max_feat_scatter(bfeat, E, mi, db, dl), orc.max_feat_scatter(f, omi, bgs, lens, E))
    perm = rng.permutation(pn).astype(np.int32)
    same(N.permutate_feature(bfeat, dev(perm)), orc.row_gather(f, perm))


@pytest.mark.parametrize("m,ofn", [(26, 8), (26, 32), (5, 3), (7, 16)])
def test_bf16_location_weight_family(N, m, ofn):
    rng = np.random.default_rng(m * ofn)
    pn = 400
    _, lens, bgs, _ = random_csr(rng, pn, 14, 0)
    E = int(lens.sum())
    lw = bf(rng.uniform(0, 1, (E, m)))
    tf = bf(rng.uniform(-1, 1, (E, m, ofn)))
    d = bf(rng.uniform(-1, 1, (pn, m, ofn)))
    dm = bf(rng.uniform(-1, 1, (pn, m)))
    blw, btf, bd, bdm, dl, db = to_bf(lw), to_bf(tf), to_bf(d), to_bf(dm), dev(lens), dev(bgs)
    close(N.location_weight_sum(blw, dl, db), orc.segment_sum(lw, lens, bgs), 1e-2)
    same(N.location_weight_sum_backward(blw, bdm, dl, db), orc.locw_sum_backward(dm, lens, bgs, E))
    close(N.location_weight_feat_sum(blw, btf, dl, db), orc.locw_feat_sum(lw, tf, lens, bgs), 1e-2)
    gdlw, gdt = N.location_weight_feat_sum_backward(blw, btf, bd, dl, db)
    odlw, odt = orc.locw_feat_sum_backward(lw, tf, d, lens, bgs)
    same(gdt, bf(odt))  # one fp32 product, one rounding
    close(gdlw, odlw, 1e-2)


def test_bf16_gradient_wiring(N):
    """the autograd Functions of neighbor_ops carry bf16 through forward and backward unchanged in dtype"""
    from pointcloudsegmentation_b200 import neighbor_ops as NO

    rng = np.random.default_rng(0)
    pn, C = 800, 16
    nidxs, lens, bgs, cens = random_csr(rng, pn, 10, 1)
    x = to_bf(bf(rng.uniform(-1, 1, (pn, C)))).requires_grad_(True)
    dn, dl, db, dc = dev(nidxs), dev(lens), dev(bgs), dev(cens)
    s = NO.neighbor_scatter(x, dn, dl, db, use_diff=True)
    p, _ = NO.neighbor_max_feat_gather(s, dl, db)
    q = NO.neighbor_sum_feat_gather(s, dc, dl, db)
    (p.float().sum() + q.float().sum()).backward()
    assert x.grad.dtype == torch.bfloat16 and torch.isfinite(x.grad.float()).all()


def test_mixed_dtypes_are_rejected(N):
    lw = torch.zeros(4, 2, device="cuda", dtype=torch.bfloat16)
    tf = torch.zeros(4, 2, 8, device="cuda")
    l = torch.tensor([4], dtype=torch.int32, device="cuda"); b = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError):
        N.location_weight_feat_sum(lw, tf, l, b)
    with pytest.raises(ValueError):
        N.neighbor_sum_feat_gather(torch.zeros(4, 2, device="cuda", dtype=torch.float16),
                                   torch.zeros(4, dtype=torch.int32, device="cuda"), l, b)

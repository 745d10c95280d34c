"""GPU: fused_ops.dense_mlp3 (csrc/dense_mlp.cu) against a plain PyTorch fp32/fp64 reference of the same
dense-concat MLP, forward and all gradients.  Tolerance 1e-5 relative (fp32 FMA chains of <= 112 terms)."""
from __future__ import annotations

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def reference(u, w2, w3, wo):
    B, A = w2.shape
    C = w3.shape[0]
    u1, u2, u3, uo = u.split([A, B, C, wo.shape[0]], 1)
    h1 = F.relu(u1)
    h2 = F.relu(u2 + h1 @ w2.t())
    h3 = F.relu(u3 + torch.cat([h2, h1], 1) @ w3.t())
    return uo + torch.cat([h3, h2, h1], 1) @ wo.t()


@pytest.mark.parametrize("dims", [(8, 8, 16, 32), (16, 16, 32, 64), (16, 16, 16, 48)])
@pytest.mark.parametrize("rows", [1, 777, 70001])
def test_dense_mlp3(dims, rows):
    from pointcloudsegmentation_b200.fused_ops import dense_mlp3, dense_mlp3_supported

    A, B, C, O = dims
    assert dense_mlp3_supported(*dims) and not dense_mlp3_supported(7, 8, 16, 32)
    g = torch.Generator(device="cuda").manual_seed(rows + A)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    u, w2, w3, wo, dy = r(rows, A + B + C + O), r(B, A) / 3, r(C, A + B) / 4, r(O, A + B + C) / 6, r(rows, O)
    args = [t.clone().requires_grad_(True) for t in (u, w2, w3, wo)]
    y = dense_mlp3(*args)
    y.backward(dy)
    ref_args = [t.double().clone().requires_grad_(True) for t in (u, w2, w3, wo)]
    yr = reference(*ref_args)
    yr.backward(dy.double())
    assert torch.allclose(y.double(), yr, rtol=1e-5, atol=1e-5)
    for a, b, name in zip(args, ref_args, ("du", "dw2", "dw3", "dwo")):
        scale = max(1.0, float(b.grad.abs().max()))
        err = float((a.grad.double() - b.grad).abs().max()) / scale
        assert err <= (1e-5 if name == "du" else 2e-5), (name, err)


@pytest.mark.parametrize("H", [8, 64, 96, 128, 256])
@pytest.mark.parametrize("pn", [1, 257, 6000])
def test_edge_combine(H, pn):
    """u[e] = pa[centre] + pb[neighbour] + wx sxyz[e], and its three gradients, vs stock PyTorch in float64."""
    import numpy as np

    from conftest import random_csr
    from pointcloudsegmentation_b200.fused_ops import edge_combine

    rng = np.random.default_rng(H * 7 + pn)
    nidxs, lens, bgs, cens = (torch.from_numpy(a).cuda() for a in random_csr(rng, pn, 9))
    E = nidxs.shape[0]
    g = torch.Generator(device="cuda").manual_seed(pn + H)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    pa, pb, sx, wx, du = r(pn, H), r(pn, H), r(E, 3), r(H, 3), r(E, H)
    args = [t.clone().requires_grad_(True) for t in (torch.cat([pa, pb], 1), wx)]
    u = edge_combine(args[0], sx, args[1], nidxs, lens, bgs, cens)
    u.backward(du)
    ref = [t.double().clone().requires_grad_(True) for t in (torch.cat([pa, pb], 1), wx)]
    ur = ref[0][cens.long(), :H] + ref[0][nidxs.long(), H:] + sx.double() @ ref[1].t()
    ur.backward(du.double())
    assert torch.allclose(u.double(), ur, rtol=1e-6, atol=1e-6)
    for a, b, name in zip(args, ref, ("dpab", "dwx")):
        scale = max(1.0, float(b.grad.abs().max()))
        assert float((a.grad.double() - b.grad).abs().max()) / scale <= 1e-5, name


def test_edge_combine_rejects_bad_width():
    from pointcloudsegmentation_b200._lib import PcsError
    from pointcloudsegmentation_b200.fused_ops import edge_combine

    z = torch.zeros(4, 12, device="cuda")
    i = torch.zeros(4, dtype=torch.int32, device="cuda")
    with pytest.raises(PcsError):
        edge_combine(z, torch.zeros(4, 3, device="cuda"), torch.zeros(6, 3, device="cuda"), i,
                     torch.ones(4, dtype=torch.int32, device="cuda"), torch.arange(4, dtype=torch.int32, device="cuda"), i)


@pytest.mark.parametrize("dims", [(8, 8, 16, 32), (16, 16, 32, 64), (16, 16, 16, 48)])
@pytest.mark.parametrize("pn", [1, 300, 9001])
def test_dense_mlp3_pool(dims, pn):
    """MLP + segmented max as one differentiable op == dense_mlp3 followed by the a5/a6 ops: pooled values
    and arg-max bit-identical, du bit-identical (same kernel arithmetic), weight gradients to 2e-5
    (the pooled dWo sums pn*O terms instead of csum*O, the rest of which are exact zeros)."""
    import numpy as np

    from conftest import random_csr
    from pointcloudsegmentation_b200 import neighbor_ops
    from pointcloudsegmentation_b200.fused_ops import dense_mlp3, dense_mlp3_pool

    A, B, C, O = dims
    rng = np.random.default_rng(pn + O)
    _, lens, bgs, cens = (torch.from_numpy(a).cuda() for a in random_csr(rng, pn, 14))  # includes empty rows
    E = cens.shape[0]
    g = torch.Generator(device="cuda").manual_seed(pn + A)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    u, w2, w3, wo, dp = r(E, A + B + C + O), r(B, A) / 3, r(C, A + B) / 4, r(O, A + B + C) / 6, r(pn, O)
    a1 = [t.clone().requires_grad_(True) for t in (u, w2, w3, wo)]
    pooled, midx = dense_mlp3_pool(*a1, lens, bgs, cens)
    pooled.backward(dp)
    a2 = [t.clone().requires_grad_(True) for t in (u, w2, w3, wo)]
    ref, ref_idx = neighbor_ops.neighbor_max_feat_gather(dense_mlp3(*a2), lens, bgs)
    ref.backward(dp)
    assert torch.equal(pooled, ref) and torch.equal(midx, ref_idx)
    assert torch.equal(a1[0].grad, a2[0].grad)
    for x, y, name in zip(a1[1:], a2[1:], ("dw2", "dw3", "dwo")):
        scale = max(1.0, float(y.grad.abs().max()))
        assert float((x.grad - y.grad).abs().max()) / scale <= 2e-5, name


@pytest.mark.parametrize("OJ", [(32, 32), (48, 48), (64, 64)])
@pytest.mark.parametrize("max_len", [3, 40, 150])
def test_pool_linear_wgrad_long_and_empty_rows(OJ, max_len):
    """pcs_pool_linear_wgrad (rows staged 32 at a time in shared memory): centres with 0, 1, exactly 32, 33 and up to 150
    rows, against  dWo[f,:] = sum_i dpool[i,f] * x[bgs[i] + max_idxs[i,f], :]  in float64, and against the round-1 gather
    kernel (pool_wgrad_staged = 0)."""
    import numpy as np

    from pointcloudsegmentation_b200._lib import check, lib
    from pointcloudsegmentation_b200.fused_ops import _stream, column_sum

    O, J = OJ
    rng = np.random.default_rng(O + max_len)
    pn = 3000
    lens = rng.integers(0, max_len + 1, pn).astype(np.int32)
    lens[:6] = [0, 1, 32, 33, max_len, 0]
    begs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int32)
    cs = int(lens.sum())
    midx = (rng.random((pn, O)) * np.maximum(lens, 1)[:, None]).astype(np.int32)
    g = torch.Generator(device="cuda").manual_seed(max_len)
    x = torch.randn(cs, J, device="cuda", generator=g)
    dpool = torch.randn(pn, O, device="cuda", generator=g)
    L, Bg, Mi = (torch.from_numpy(a).cuda() for a in (lens, begs, midx))
    live = (L > 0).double()[:, None]
    rows = (Bg.long()[:, None] + Mi.long()).clamp(max=cs - 1)                    # [pn, O]
    ref = torch.einsum("if,ifj->fj", dpool.double() * live, x.double()[rows])    # [O, J]
    outs = []
    try:
        for mode in (1, 0):
            assert lib.pcs_set_option(b"pool_wgrad_staged", mode) == 0
            P = int(lib.pcs_pool_linear_wgrad_partials(pn, O))
            part = torch.empty((P, O, J), device="cuda")
            check(lib.pcs_pool_linear_wgrad(dpool.data_ptr(), Mi.data_ptr(), L.data_ptr(), Bg.data_ptr(), x.data_ptr(), pn, cs, O, J,
                                            part.data_ptr(), _stream()), "pool_linear_wgrad")
            outs.append(column_sum(part))
    finally:
        lib.pcs_set_option(b"pool_wgrad_staged", 1)
    scale = max(1.0, float(ref.abs().max()))
    for o in outs:
        assert float((o.double() - ref).abs().max()) / scale <= 1e-5


def test_set_option_and_ffma_fallback_agree_with_tensor_core_path():
    """pcs_set_option: unknown names fail; mlp_tensor_cores=0 selects the fp32 FFMA kernels, which must agree
    with the default tensor-core (3xTF32) kernels to 1e-5 — forward, du and all weight gradients."""
    from conftest import random_csr
    import numpy as np

    from pointcloudsegmentation_b200._lib import lib
    from pointcloudsegmentation_b200.fused_ops import dense_mlp3_pool

    assert lib.pcs_set_option(b"no_such_option", 1) != 0
    A, B, C, O = 8, 8, 16, 32
    rng = np.random.default_rng(5)
    _, lens, bgs, cens = (torch.from_numpy(a).cuda() for a in random_csr(rng, 4000, 14))
    E = cens.shape[0]
    g = torch.Generator(device="cuda").manual_seed(5)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    u, w2, w3, wo, dp = r(E, A + B + C + O), r(B, A) / 3, r(C, A + B) / 4, r(O, A + B + C) / 6, r(4000, O)
    res = []
    try:
        for mode in (1, 0):
            assert lib.pcs_set_option(b"mlp_tensor_cores", mode) == 0
            a = [t.clone().requires_grad_(True) for t in (u, w2, w3, wo)]
            pooled, midx = dense_mlp3_pool(*a, lens, bgs, cens)
            pooled.backward(dp)
            res.append((pooled.detach(), [t.grad for t in a]))
    finally:
        lib.pcs_set_option(b"mlp_tensor_cores", 1)
    (p1, g1), (p0, g0) = res
    assert torch.allclose(p1, p0, rtol=1e-5, atol=1e-5)
    # du: a near-tie (two rows within ~1e-6 of each other in one channel) may pick a different arg-max row in
    # the two arithmetic paths and move that gradient entry to the other row — allow a handful of such rows
    bad = ((g1[0] - g0[0]).abs() > 1e-5 * max(1.0, float(g0[0].abs().max()))).any(1)
    assert int(bad.sum()) <= 8, int(bad.sum())
    for x, y in zip(g1[1:], g0[1:]):  # weight gradients: a flipped row changes one term out of ~10^5
        scale = max(1.0, float(y.abs().max()))
        assert float((x - y).abs().max()) / scale <= 1e-4


@pytest.mark.parametrize("shape", [(0, 8), (1, 5), (31, 128), (33, 35), (70001, 64), (4097, 131), (592, 32, 24),
                                   (2000, 1030), (300, 16, 3), (100000, 13)])
def test_column_sum(shape):
    """out[c] = sum_r x[r, c] (csrc/column_sum.cu) vs float64; <= 1e-5 of the largest column sum of |x|;
    twice the same bits (fixed summation order)."""
    from pointcloudsegmentation_b200 import fused_ops

    column_sum = lambda t: fused_ops.column_sum(t, small_stock=False)  # noqa: E731  (always the CUDA kernel)
    g = torch.Generator(device="cuda").manual_seed(sum(shape))
    x = torch.randn(*shape, device="cuda", generator=g)
    y = column_sum(x)
    assert y.shape == x.shape[1:]
    ref = x.double().sum(0)
    scale = max(1.0, float(x.double().abs().sum(0).max())) if shape[0] else 1.0
    assert float((y.double() - ref).abs().max()) <= 1e-5 * scale
    assert torch.equal(y, column_sum(x))
    # a view whose rows start 4 bytes off a 16-byte boundary (vector width falls back)
    if shape[0] > 1 and len(shape) == 2:
        flat = torch.randn(shape[0] * shape[1] + 1, device="cuda", generator=g)
        v = flat[1:].view(shape)
        assert float((column_sum(v).double() - v.double().sum(0)).abs().max()) <= 1e-5 * max(1.0, float(v.abs().sum(0).max()))

"""GPU: fused_ops.linear_tc (csrc/linear_tc.cu, tcgen05 3xTF32) against a float64 reference of the same
product.  Tolerance: 2e-5 of the result's max-norm (stated in csrc/linear_tc.cu: the split drops the lo*lo
term and the tensor core accumulates in fp32; the stock fp32 GEMM is at ~1e-6 on the same inputs)."""
from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-5


def _check(y, ref):
    scale = max(1.0, float(ref.abs().max()))
    err = float((y.double() - ref).abs().max())
    assert err <= TOL * scale, f"max err {err:.3e} vs scale {scale:.3e}"


@pytest.mark.parametrize("shape", [(1, 8, 16), (129, 8, 40), (255, 33, 17), (257, 32, 256), (1000, 643, 256),
                                   (5000, 131, 512), (4097, 256, 643), (70001, 35, 128), (3000, 64, 134),
                                   (2048, 1474, 512), (3000, 64, 1300)])  # (the last: bias wider than its shared-memory copy)
@pytest.mark.parametrize("mode", ["plain", "bias", "bias_add_relu"])
def test_linear_tc_matches_float64(shape, mode):
    from pointcloudsegmentation_b200.fused_ops import linear_tc

    M, K, N = shape
    g = torch.Generator(device="cuda").manual_seed(M + K + N)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    b = torch.randn(N, device="cuda", generator=g) if mode != "plain" else None
    add = torch.randn(M, N, device="cuda", generator=g) if mode == "bias_add_relu" else None
    y = linear_tc(x, w, b, add, relu=(mode == "bias_add_relu"))
    ref = x.double() @ w.double().t()
    if b is not None:
        ref = ref + b.double()
    if add is not None:
        ref = torch.relu(ref + add.double())
    _check(y, ref)


def test_linear_tc_strided_operands():
    """Column slices of wider tensors (x, addend, out) and transposed / sliced weights are passed as they are."""
    from pointcloudsegmentation_b200.fused_ops import linear_tc

    g = torch.Generator(device="cuda").manual_seed(7)
    M = 3001
    xw = torch.randn(M, 200, device="cuda", generator=g)
    x = xw[:, 5:104]  # K = 99, row stride 200, unaligned start
    W = torch.randn(300, 160, device="cuda", generator=g)
    w_t = W[10:109, 3:150].t()  # [N=147, K=99] view with strides (1, 160)
    addw = torch.randn(M, 400, device="cuda", generator=g)
    add = addw[:, 1:148]
    outw = torch.zeros(M, 151, device="cuda")
    y = linear_tc(x, w_t, None, add, out=outw[:, 2:149])
    ref = x.double() @ w_t.double().t() + add.double()
    _check(y, ref)
    assert float(outw[:, :2].abs().max()) == 0.0 and float(outw[:, 149:].abs().max()) == 0.0  # neighbours untouched
    # dX form: g [M,N] @ w [N,K]  ==  linear_tc(g, w.t())
    gy = torch.randn(M, 147, device="cuda", generator=g)
    w = torch.randn(147, 99, device="cuda", generator=g)
    _check(linear_tc(gy, w.t()), gy.double() @ w.double())


@pytest.mark.parametrize("N", [5, 33, 99, 130, 260])
@pytest.mark.parametrize("ld_mod", [0, 1, 2, 3])
def test_linear_tc_every_output_alignment(N, ld_mod):
    """The epilogue writes 16-byte segments wherever they fall inside a row piece: every phase of the output's first
    element and every row stride modulo 4, addend in phase with the output and out of phase, bias unaligned; the
    columns around the output slice must stay untouched."""
    from pointcloudsegmentation_b200.fused_ops import linear_tc

    M, K = 300, 40
    g = torch.Generator(device="cuda").manual_seed(N * 4 + ld_mod)
    x = torch.randn(M, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    bias = torch.randn(N + 4, device="cuda", generator=g)[(4 if ld_mod == 0 else 1):][:N]  # 16-byte aligned only for ld_mod 0
    ld = N + 8
    ld += (ld_mod - ld) % 4
    for off in range(4):
        for add_off in (off, (off + 1) % 4, None):
            outw = torch.full((M, ld), -7.0, device="cuda")
            out = outw[:, off:off + N]
            add = None
            if add_off is not None:
                add = torch.randn(M, ld + 4 * (off % 2), device="cuda", generator=g)[:, add_off:add_off + N]
            y = linear_tc(x, w, bias, add, relu=add is not None, out=out)
            assert y.data_ptr() == out.data_ptr()
            ref = x.double() @ w.double().t() + bias.double()
            if add is not None:
                ref = torch.relu(ref + add.double())
            _check(y, ref)
            assert bool((outw[:, :off] == -7.0).all()) and bool((outw[:, off + N:] == -7.0).all())


def test_linear_tc_rejects_bad_arguments():
    from pointcloudsegmentation_b200.fused_ops import linear_tc

    x = torch.zeros(4, 8, device="cuda")
    with pytest.raises(ValueError):
        linear_tc(x, torch.zeros(3, 9, device="cuda"))
    with pytest.raises(ValueError):
        linear_tc(x.cpu(), torch.zeros(3, 8))
    with pytest.raises(ValueError):
        linear_tc(x, torch.zeros(3, 8, device="cuda"), bias=torch.zeros(4, device="cuda"))
    assert linear_tc(torch.zeros(0, 8, device="cuda"), torch.zeros(3, 8, device="cuda")).shape == (0, 3)


def test_network_layers_route_tall_products_to_linear_tc():
    """network.linear_fwd sends tall activations through the tensor-core kernel (launch counter moves by 2:
    weight image + GEMM) and short ones to the stock GEMM."""
    import pointcloudsegmentation_b200 as pcs
    from pointcloudsegmentation_b200.network import linear_fwd

    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(9000, 40, device="cuda", generator=g)
    w = torch.randn(64, 40, device="cuda", generator=g)
    n0 = pcs.launch_count()
    y = linear_fwd(x, w)
    assert pcs.launch_count() - n0 == 2
    _check(y, x.double() @ w.double().t())
    n0 = pcs.launch_count()
    linear_fwd(x[:100], w)
    assert pcs.launch_count() == n0


@pytest.mark.parametrize("shape", [(17, 16, 8), (300, 32, 32), (4097, 13, 387), (20000, 128, 35), (9000, 256, 643),
                                   (9000, 512, 131), (70001, 64, 134), (5000, 130, 300), (2500, 1474, 40)])
@pytest.mark.parametrize("with_bias", [False, True])
def test_linear_tc_wgrad_matches_float64(shape, with_bias):
    """dW = g^T x (+ db = column sums of g) against float64; same 2e-5 max-norm bar as the forward kernel."""
    from pointcloudsegmentation_b200.fused_ops import linear_tc_wgrad

    M, N, K = shape
    gen = torch.Generator(device="cuda").manual_seed(M + N + K)
    g = torch.randn(M, N, device="cuda", generator=gen)
    x = torch.randn(M, K, device="cuda", generator=gen)
    dw, db = linear_tc_wgrad(g, x, with_bias)
    assert dw.shape == (N, K)
    _check(dw, g.double().t() @ x.double())
    if with_bias:
        _check(db, g.double().sum(0))
    else:
        assert db is None


def test_linear_tc_wgrad_strided_and_deterministic():
    from pointcloudsegmentation_b200.fused_ops import linear_tc_wgrad

    gen = torch.Generator(device="cuda").manual_seed(11)
    M = 12345
    gw = torch.randn(M, 300, device="cuda", generator=gen)
    xw = torch.randn(M, 200, device="cuda", generator=gen)
    g, x = gw[:, 7:139], xw[:, 4:103]  # N = 132, K = 99; unaligned starts, wide row strides
    dw, db = linear_tc_wgrad(g, x, True)
    _check(dw, g.double().t() @ x.double())
    _check(db, g.double().sum(0))
    dw2, db2 = linear_tc_wgrad(g, x, True)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)

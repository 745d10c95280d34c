"""GPU: graph_conv_layers (the reference's L3 layers, tf_ops/graph_conv_layer.py:74-166,573-627 and
model_pgnet.py:4-32) on the CUDA ops, against the same layer code run over a plain-PyTorch float64 indexing
namespace (`feats[nidxs] - feats[cens]`, `index_add_`).  Outputs and every gradient; tolerance 1e-5 relative
to the largest magnitude of the tensor compared (fp32 sums over <= 16 neighbours and <= 67-term dot products)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


class TorchOps:
    """The four differentiable ops the layers call, written with stock indexing (any dtype, autograd by torch)."""

    def __init__(self, cens):
        self.cens = cens.long()

    def neighbor_scatter(self, ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False):
        s = ifeats[nidxs.long()]
        return s - ifeats[self.cens] if use_diff else s

    def neighbor_sum_feat_gather(self, ifeats, cidxs, nidxs_lens, nidxs_bgs):
        out = torch.zeros(nidxs_lens.shape[0], ifeats.shape[1], dtype=ifeats.dtype, device=ifeats.device)
        return out.index_add_(0, cidxs.long(), ifeats)

    def location_weight_sum(self, lw, nidxs_lens, nidxs_bgs):
        return self.neighbor_sum_feat_gather(lw, self.cens, nidxs_lens, nidxs_bgs)

    def location_weight_feat_sum(self, lw, feats, nidxs_lens, nidxs_bgs):
        w = (lw.unsqueeze(2) * feats).reshape(feats.shape[0], -1)
        return self.neighbor_sum_feat_gather(w, self.cens, nidxs_lens, nidxs_bgs).reshape(-1, *feats.shape[1:])


def _graph(pn, seed, max_len=16):
    from conftest import random_csr

    rng = np.random.default_rng(seed)
    nidxs, lens, bgs, cens = (torch.from_numpy(a).cuda() for a in random_csr(rng, pn, max_len, min_len=1))  # radius neighbourhoods hold the centre
    return nidxs, lens, bgs, cens


def _close(a, b, name, tol=1e-5):
    a, b = a.detach(), b.detach()
    scale = max(1.0, float(b.abs().max()))
    err = float((a.double() - b).abs().max()) / scale
    assert err <= tol, (name, err)


def _check(fn, tensors, names, tol=1e-5):
    """fn(ops_or_None, *tensors) -> output; runs the CUDA fp32 path and the float64 indexing path."""
    a32 = [t.clone().requires_grad_(True) for t in tensors]
    a64 = [t.double().clone().requires_grad_(True) for t in tensors]
    y32, y64 = fn(False, *a32), fn(True, *a64)
    _close(y32, y64, "out", tol)
    g = torch.Generator(device="cuda").manual_seed(11)
    dy = torch.randn(y32.shape, device="cuda", generator=g)
    y32.backward(dy)
    y64.backward(dy.double())
    for a, b, n in zip(a32, a64, names):
        _close(a.grad, b.grad, "d" + n, tol)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("no_sum", [False, True])
@pytest.mark.parametrize("pn,ifn,m,ofn", [(1, 4, 3, 2), (300, 16, 9, 8), (2000, 13, 27, 16)])
def test_graph_conv_xyz_feats(pn, ifn, m, ofn, no_sum, fused):
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn + m)
    g = torch.Generator(device="cuda").manual_seed(pn)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    xyz, feats, pw, pmiu = r(pn, 3) * 0.3, r(pn, ifn), r(3 + ifn, m, ofn) / 4, r(3, m)

    def fn(ref, xyz, feats, pw, pmiu):
        ops = TorchOps(cens) if ref else None
        out, lw, lw_sum = L.graph_conv_xyz_feats_impl(xyz, feats, cens, nidxs, lens, bgs, pw, compute_lw=True,
                                                      pmiu=pmiu, no_sum=no_sum, fused=fused and not ref, ops=ops)
        # the reference reuses (lw, lw_sum) in the following feats-only convolutions of the stage
        out2 = L.graph_conv_feats_impl(feats, cens, nidxs, lens, bgs, pw[3:], lw, lw_sum, no_sum,
                                       fused and not ref, ops)
        out3 = L.graph_diff_conv_feats_impl(feats, cens, nidxs, lens, bgs, pw[3:], lw, lw_sum, no_sum,
                                            fused and not ref, ops)
        return torch.cat([out, out2, out3], 1)

    _check(fn, (xyz, feats, pw, pmiu), ("xyz", "feats", "pw", "pmiu"), tol=2e-5)


@pytest.mark.parametrize("pn,m,ofn", [(257, 9, 16), (1500, 27, 8)])
def test_graph_conv_xyz(pn, m, ofn):
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn * 3 + m)
    g = torch.Generator(device="cuda").manual_seed(pn)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    xyz, pw, pmiu = r(pn, 3) * 0.3, r(3, m, ofn), r(3, m)

    def fn(ref, xyz, pw, pmiu):
        ops = TorchOps(cens) if ref else None
        return L.graph_conv_xyz_impl(xyz, cens, nidxs, lens, bgs, pw, compute_lw=True, pmiu=pmiu,
                                     fused=not ref, ops=ops)[0]

    _check(fn, (xyz, pw, pmiu), ("xyz", "pw", "pmiu"), tol=2e-5)


def _module_pair(make):
    torch.manual_seed(5)
    m32 = make().cuda()
    m64 = make().cuda().double()
    m64.load_state_dict({k: v.double() for k, v in m32.state_dict().items()})
    return m32, m64


def _check_module(m32, m64, call, tensors, tol=2e-5):
    a32 = [t.clone().requires_grad_(True) for t in tensors]
    a64 = [t.double().clone().requires_grad_(True) for t in tensors]
    y32, y64 = call(m32, *a32), call(m64, *a64)
    _close(y32, y64, "out", tol)
    g = torch.Generator(device="cuda").manual_seed(3)
    dy = torch.randn(y32.shape, device="cuda", generator=g)
    y32.backward(dy)
    y64.backward(dy.double())
    for i, (a, b) in enumerate(zip(a32, a64)):
        _close(a.grad, b.grad, f"dinput{i}", tol)
    for (n, p), (_, q) in zip(m32.named_parameters(), m64.named_parameters()):
        _close(p.grad, q.grad, "d" + n, tol)


@pytest.mark.parametrize("pn,ifn,ofn", [(1, 4, 4), (400, 8, 16), (1500, 16, 12)])
def test_graph_conv_edge(pn, ifn, ofn):
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn + ifn, 12)
    g = torch.Generator(device="cuda").manual_seed(pn)
    feats, sxyz = torch.randn(pn, ifn, device="cuda", generator=g), torch.randn(nidxs.shape[0], 3, device="cuda", generator=g)
    m32, m64 = _module_pair(lambda: L.GraphConvEdge(ifn, [8, 8], ofn))
    m64.ops = TorchOps(cens)
    _check_module(m32, m64, lambda m, s, f: m(s, f, nidxs, lens, bgs, cens), (sxyz, feats))


@pytest.mark.parametrize("pn,ofn", [(300, 16), (1200, 5)])
def test_graph_conv_edge_xyz(pn, ofn):
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn + ofn, 12)
    g = torch.Generator(device="cuda").manual_seed(pn)
    sxyz = torch.randn(nidxs.shape[0], 3, device="cuda", generator=g)
    m32, m64 = _module_pair(lambda: L.GraphConvEdgeXyz([8, 8], ofn))
    m64.ops = TorchOps(cens)
    _check_module(m32, m64, lambda m, s: m(s, nidxs, lens, bgs, cens), (sxyz,))


@pytest.mark.parametrize("pn,dim,out", [(1, 8, 8), (500, 16, 32), (1800, 35, 24)])
def test_ecd_feats(pn, dim, out):
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn + dim, 12)
    g = torch.Generator(device="cuda").manual_seed(pn)
    feats, ec = torch.randn(pn, dim, device="cuda", generator=g), torch.randn(nidxs.shape[0], 3, device="cuda", generator=g)
    m32, m64 = _module_pair(lambda: L.EcdFeats(dim, [8, 8], [16], out))
    m64.ops = TorchOps(cens)
    _check_module(m32, m64, lambda m, e, f: m(e, f, nidxs, lens, bgs, cens), (ec, feats))


@pytest.mark.parametrize("pn,dim,out", [(1, 8, 8), (500, 16, 32), (1800, 32, 24)])
def test_ecd_feats_split_weights(pn, dim, out):
    """The split-weight evaluation of the phi-MLP + gate (one point-level GEMM + fused_ops.edge_combine instead of the
    difference tensor and the dense concats; taken when the edge coordinates need no gradient, as in the towers) against
    the literal float64 evaluation: output, d(feats) and every parameter gradient."""
    from pointcloudsegmentation_b200 import graph_conv_layers as L

    nidxs, lens, bgs, cens = _graph(pn, pn + dim, 12)
    g = torch.Generator(device="cuda").manual_seed(pn)
    feats, ec = torch.randn(pn, dim, device="cuda", generator=g), torch.randn(nidxs.shape[0], 3, device="cuda", generator=g)
    m32, m64 = _module_pair(lambda: L.EcdFeats(dim, [dim // 2, dim // 2], [dim // 2, dim // 2], out))
    m64.ops = TorchOps(cens)
    taken = []
    orig = m32._gate_split
    m32._gate_split = lambda *a: (taken.append(1), orig(*a))[1]
    _check_module(m32, m64, lambda m, f: m(ec if m is m32 else ec.double(), f, nidxs, lens, bgs, cens), (feats,), tol=2e-5)
    assert taken, "the split path was not exercised"

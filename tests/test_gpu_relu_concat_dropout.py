"""GPU: relu + concat + dropout in one pass (csrc/relu_concat_dropout.cu, classifier_v3 model.py:783-793) against
the numpy restatement (oracle/dropout_mask.py), bit for bit — values, mask and both gradients."""
import numpy as np
import pytest
import torch

from oracle import dropout_mask as dm

pytestmark = pytest.mark.gpu


def _ops():
    from pointcloudsegmentation_b200 import fused_ops

    return fused_ops


def _rand(M, Na, Nb, seed=0):
    rng = np.random.default_rng(seed)
    return rng.standard_normal((M, Na)).astype(np.float32), rng.standard_normal((M, Nb)).astype(np.float32)


@pytest.mark.parametrize("M,Na,Nb", [(1000, 512, 131), (333, 256, 131), (77, 5, 0), (50, 3, 2), (1, 1, 0), (129, 31, 1), (4096, 128, 0)])
@pytest.mark.parametrize("keep", [1.0, 0.7, 0.5])
def test_forward_backward_bit_exact(M, Na, Nb, keep):
    F = _ops()
    a, b = _rand(M, Na, Nb, M + Na)
    seed = 0x1234_5678_9ABC_DEF0 ^ (M * 7919)
    ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    out = F.relu_concat_dropout_forward(ta, tb, keep, seed)
    ref = dm.relu_concat_dropout_forward(a, b, keep, seed)
    assert np.array_equal(out.cpu().numpy(), ref)
    g = np.random.default_rng(1).standard_normal(ref.shape).astype(np.float32)
    da, db = F.relu_concat_dropout_backward(torch.from_numpy(g).cuda(), out, Na, keep, seed)
    rda, rdb = dm.relu_concat_dropout_backward(g, a, keep, seed)
    assert np.array_equal(da.cpu().numpy(), rda)
    if Nb:
        assert np.array_equal(db.cpu().numpy(), rdb)
    else:
        assert db is None


def test_strided_inputs_and_empty():
    F = _ops()
    big = torch.randn(500, 300, device="cuda")
    a, b = big[:, 3:200], big[:, 205:260]  # column slices: row stride 300, unaligned starts
    out = F.relu_concat_dropout_forward(a, b, 0.7, 99)
    ref = dm.relu_concat_dropout_forward(a.cpu().numpy(), b.cpu().numpy(), 0.7, 99)
    assert np.array_equal(out.cpu().numpy(), ref)
    e = F.relu_concat_dropout_forward(torch.empty(0, 8, device="cuda"), torch.empty(0, 3, device="cuda"), 0.7, 1)
    assert e.shape == (0, 11)


def test_autograd_matches_stock_ops_without_dropout():
    F = _ops()
    a = torch.randn(3000, 64, device="cuda", requires_grad=True)
    b = torch.randn(3000, 17, device="cuda", requires_grad=True)
    w = torch.randn(3000, 81, device="cuda")
    for training, keep in ((False, 0.7), (True, 1.0)):
        (F.relu_concat_dropout(a, b, keep, training) * w).sum().backward()
        ga, gb = a.grad.clone(), b.grad.clone()
        a.grad = b.grad = None
        (torch.cat([torch.relu(a), b], 1) * w).sum().backward()
        assert torch.equal(ga, a.grad) and torch.equal(gb, b.grad)
        a.grad = b.grad = None


def test_autograd_with_dropout_is_consistent_and_seeded():
    F = _ops()
    a = torch.randn(2000, 40, device="cuda", requires_grad=True)
    b = torch.randn(2000, 9, device="cuda", requires_grad=True)
    torch.manual_seed(7)
    y1 = F.relu_concat_dropout(a, b, 0.7, True)
    torch.manual_seed(7)
    y2 = F.relu_concat_dropout(a, b, 0.7, True)
    y3 = F.relu_concat_dropout(a, b, 0.7, True)
    assert torch.equal(y1, y2) and not torch.equal(y1, y3)
    x = torch.cat([torch.relu(a), b], 1).detach()
    kept = y1 != 0
    assert torch.allclose(y1[kept], x[kept] / 0.7, rtol=1e-6, atol=0)
    y1.sum().backward()
    # d/da = 1/keep where the output is positive; d/db = 1/keep where kept (b != 0 almost surely, so kept <=> y != 0)
    assert torch.allclose(a.grad, (y1[:, :40] > 0).float() / 0.7, rtol=1e-6, atol=0)
    assert torch.allclose(b.grad, (y1[:, 40:] != 0).float() / 0.7, rtol=1e-6, atol=0)


def test_full_size_properties():
    """One level-0 classifier tensor of a 16-block step: kept fraction, value rule — no CPU oracle needed."""
    F = _ops()
    M, Na, Nb, keep = 133399, 512, 131, 0.7
    a = torch.randn(M, Na, device="cuda")
    b = torch.randn(M, Nb, device="cuda")
    y = F.relu_concat_dropout_forward(a, b, keep, 2026)
    x = torch.cat([torch.relu(a), b], 1)
    nz = x != 0
    frac = float(((y != 0) & nz).sum()) / float(nz.sum())
    assert abs(frac - keep) < 5 * np.sqrt(keep * (1 - keep) / float(nz.sum()))
    scale = torch.tensor(1.0, dtype=torch.float32, device="cuda") / torch.tensor(keep, dtype=torch.float32, device="cuda")
    assert bool(((y == 0) | (y == x * scale)).all())
    col_frac = ((y != 0) & nz).float().sum(0) / nz.float().sum(0).clamp(min=1)
    assert float((col_frac - keep).abs().max()) < 0.02  # no column is favoured by the 128-element chunking


def test_bad_arguments():
    F = _ops()
    a = torch.randn(10, 4, device="cuda")
    with pytest.raises(ValueError):
        F.relu_concat_dropout_forward(a.cpu(), a, 0.7, 1)
    with pytest.raises(ValueError):
        F.relu_concat_dropout_forward(a, torch.randn(9, 4, device="cuda"), 0.7, 1)
    with pytest.raises(RuntimeError):
        F.relu_concat_dropout_forward(a, a, 0.0, 1)


def test_classifier_routes_through_the_fused_kernel():
    from pointcloudsegmentation_b200 import _lib, network

    net = network.Pointnet13DilatedEmbed(13).cuda()
    z0 = torch.randn(9000, 512, device="cuda")
    pf = torch.randn(9000, 131, device="cuda")
    net.eval()
    with torch.no_grad():
        net.classify(z0, pf)  # materialises the lazy layers through the stock path
        n0 = _lib.lib.pcs_launch_count()
        fused = net.classify(z0, pf)
        assert _lib.lib.pcs_launch_count() - n0 >= 2
        L = net.L
        net.L = type(L)(**{k: v for k, v in vars(L).items() if k != "relu_concat_dropout"})
        stock = net.classify(z0, pf)
        net.L = L
    assert torch.allclose(fused, stock, rtol=2e-5, atol=2e-5)

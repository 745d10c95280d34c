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

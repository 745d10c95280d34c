"""GPU: the S3DIS tower end to end (a21 composites + every op's gradient wiring at once): CUDA ops vs
the same network on the CPU oracle layers, same weights, same block.  Index outputs of the pooling
pipeline bit-exact; loss and weight gradients within 1e-4 relative (fp32 GEMMs on two different
devices accumulate differently; the custom-op sums themselves are held to 1e-5 in the op tests)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import cpu_layers
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed
from pointcloudsegmentation_b200.synthetic import s3dis_batch

pytestmark = pytest.mark.gpu


def build_pair():
    torch.manual_seed(0)
    cpu = Pointnet13DilatedEmbed(L=cpu_layers.layers(), split=False)  # literal reference order
    cpu.eval()
    d = s3dis_batch(3, 1, 900, 1100)
    with torch.no_grad():
        cpu(*(torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels")))
    gpu = Pointnet13DilatedEmbed().cuda()
    gpu.eval()
    with torch.no_grad():
        gpu(*(torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels")))
    gpu.load_state_dict(cpu.state_dict())
    return cpu, gpu


def test_pooling_pipeline_matches_oracle():
    from pointcloudsegmentation_b200.graph_layers import points_pooling_two_layers

    d = s3dis_batch(5, 3, 1500, 2500)
    args = [torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels")]
    bids = torch.from_numpy(d["block_ids"])
    want = cpu_layers.points_pooling_two_layers(*args, 0.15, 0.45, 3.0, block_ids=bids)
    got = points_pooling_two_layers(*(a.cuda() for a in args), 0.15, 0.45, 3.0, block_ids=bids.cuda())
    flat = lambda o: [t for x in o for t in (x if isinstance(x, (list, tuple)) else [x])]  # noqa: E731
    for g, w in zip(flat(got), flat(want)):
        if w.dtype.is_floating_point:
            # centroids: fp32 segment sums / counts (oracle sums in float64 then rounds)
            assert torch.allclose(g.cpu(), w, rtol=1e-5, atol=1e-6)
        else:
            assert torch.equal(g.cpu(), w)


def test_network_loss_and_gradients_match_cpu_oracle_network():
    cpu, gpu = build_pair()
    d = s3dis_batch(7, 2, 1200, 1600)
    a = [torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels", "block_ids")]
    lc, logits_c, lab_c = cpu(*a, 2)
    lg, logits_g, lab_g = gpu(*(t.cuda() for t in a), 2)
    assert torch.equal(lab_g.cpu(), lab_c)
    assert abs(float(lg) - float(lc)) <= 1e-4 * max(1.0, abs(float(lc)))
    assert torch.allclose(logits_g.cpu(), logits_c, rtol=1e-3, atol=1e-3)
    lc.backward()
    lg.backward()
    # Frobenius-relative, not max: a pre-activation within 1e-7 of zero can flip its ReLU mask between
    # two fp32 evaluation orders, which moves single gradient rows by ~1e-3 of the max entry.
    worst = 0.0
    for pc, pg in zip(cpu.parameters(), gpu.parameters()):
        assert (pc.grad is None) == (pg.grad is None)
        if pc.grad is None:
            continue
        scale = max(float(pc.grad.norm()), 1e-9)
        worst = max(worst, float((pg.grad.cpu() - pc.grad).norm()) / scale)
    assert worst <= 2e-3, worst


def test_batched_blocks_equal_per_block_runs():
    _, gpu = build_pair()
    d = s3dis_batch(11, 3, 1000, 1400)
    x, f, l, b = (torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids"))
    with torch.no_grad():
        _, logits, lab = gpu(x, f, l, b, 3)
        off = 0
        for k in range(3):
            m = b == k
            _, lg1, lab1 = gpu(x[m], f[m], l[m])
            n = lg1.shape[0]
            assert torch.equal(lab[off:off + n], lab1)
            assert torch.allclose(logits[off:off + n], lg1, rtol=1e-4, atol=1e-4)
            off += n


def test_prepare_ahead_thread_equals_inline_prepare():
    """pipeline.PrepareAhead (own stream, own host thread, pinned host batch -> H2D -> model.prepare) hands
    the training stream exactly what an inline prepare builds: same indices, same loss bits, for several
    batches in flight one after another while the main stream is busy."""
    from pointcloudsegmentation_b200.pipeline import PrepareAhead

    _, gpu = build_pair()
    gpu.train()  # the transpose CSRs are part of prepare in training mode
    B = 3
    host = []
    for seed in (21, 22, 23):
        d = s3dis_batch(seed, B, 900, 1300)
        host.append({k: torch.from_numpy(d[k]).pin_memory() for k in ("xyzs", "feats", "labels", "block_ids")})
    ahead = PrepareAhead(gpu, B)
    futs = [ahead.submit(lambda h=h: {k: v.cuda(non_blocking=True) for k, v in h.items()}) for h in host]
    busy = torch.randn(4096, 4096, device="cuda")
    for h, fut in zip(host, futs):
        busy = busy @ busy * 1e-4  # keep the training stream occupied while the look-ahead stream works
        prep = ahead.take(fut)
        ref = gpu.prepare(*(h[k].cuda() for k in ("xyzs", "feats", "labels", "block_ids")), B)
        for a, b in zip(_flat(prep), _flat(ref)):
            assert torch.equal(a, b)
        torch.manual_seed(1)
        la = gpu(prepared=prep)[0]
        torch.manual_seed(1)
        lb = gpu(prepared=ref)[0]
        assert torch.equal(la, lb)
        la.backward()
    ahead.close()


def _flat(x):
    if torch.is_tensor(x):
        yield x
    elif isinstance(x, (list, tuple)):
        for y in x:
            yield from _flat(y)


def test_network_matches_the_reference_cuda_kernels_end_to_end():
    """The whole tower on this repo's kernels vs the SAME tower on the reference's own unmodified CUDA kernels
    (oracle/_ref/libref.so behind oracle/libref_layers.py), same weights, one block (the reference's unit of work):
    pooling indices and permuted labels bit-exact, loss 1e-4, logits 1e-3, weight gradients 2e-3 Frobenius-relative
    (eval mode: no dropout; the two arms order their fp32 GEMM sums differently)."""
    from oracle import libref, libref_layers

    if not libref.available():
        pytest.skip("oracle/_ref/libref.so not built")
    d = s3dis_batch(21, 1, 2000, 2400)
    a = [torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels")]
    torch.manual_seed(0)
    ref = Pointnet13DilatedEmbed(L=libref_layers.layers(), split=False).cuda()
    ref.eval()
    with torch.no_grad():
        ref(*a)
    ours = Pointnet13DilatedEmbed().cuda()
    ours.eval()
    with torch.no_grad():
        ours(*a)
    ours.load_state_dict(ref.state_dict())
    pr, po = ref.prepare(*a), ours.prepare(*a)
    for tr, to in zip(pr[0] + pr[1], po[0] + po[1]):  # xyzs of the three levels, dxyzs
        assert torch.allclose(tr, to, rtol=1e-5, atol=1e-6)
    assert torch.equal(pr[3], po[3])  # permuted labels
    for lr, lo in zip(pr[4] + pr[5] + pr[6], po[4] + po[5] + po[6]):  # vlens, vbegs, vcens
        assert torch.equal(lr, lo)
    for gr, go in zip(pr[7], po[7]):  # the nine graphs: nidxs, nlens, nbegs, ncens bit-identical
        for k in range(1, 5):
            assert torch.equal(gr[k], go[k])
    lr, logits_r, lab_r = ref(*a)
    lo, logits_o, lab_o = ours(*a)
    assert torch.equal(lab_r, lab_o)
    assert abs(float(lo.detach()) - float(lr.detach())) <= 1e-4 * max(1.0, abs(float(lr.detach())))
    assert torch.allclose(logits_o, logits_r, rtol=1e-3, atol=1e-3)
    lr.backward()
    lo.backward()
    worst = 0.0
    for p_r, p_o in zip(ref.parameters(), ours.parameters()):
        assert (p_r.grad is None) == (p_o.grad is None)
        if p_r.grad is None:
            continue
        worst = max(worst, float((p_o.grad - p_r.grad).norm()) / max(float(p_r.grad.norm()), 1e-9))
    assert worst <= 2e-3, worst


def test_semantic3d_dense_model_matches_cpu_oracle_network():
    """Semantic3DDense (BASELINE configs[4]) on the CUDA ops vs the same network on the CPU oracle layers: same
    weights, same dense crop; loss 1e-4, logits 1e-3, weight gradients 2e-3 Frobenius-relative."""
    from pointcloudsegmentation_b200.network import Semantic3DDense
    from test_network_shapes import _semantic3d_inputs

    a = [torch.from_numpy(v) for v in _semantic3d_inputs()]
    torch.manual_seed(0)
    cpu = Semantic3DDense(L=cpu_layers.layers(), split=False).eval()
    with torch.no_grad():
        cpu(*a)
    gpu = Semantic3DDense().cuda().eval()
    g = [t.cuda() for t in a]
    with torch.no_grad():
        gpu(*g)
    gpu.load_state_dict(cpu.state_dict())
    lc, logits_c, tgt_c, unk_c = cpu(*a)
    lg, logits_g, tgt_g, unk_g = gpu(*g)
    assert torch.equal(tgt_g.cpu(), tgt_c) and torch.equal(unk_g.cpu(), unk_c)
    assert abs(float(lg.detach()) - float(lc.detach())) <= 1e-4 * max(1.0, abs(float(lc.detach())))
    assert torch.allclose(logits_g.cpu(), logits_c, rtol=1e-3, atol=1e-3)
    lc.backward()
    lg.backward()
    worst = 0.0
    for pc, pg in zip(cpu.parameters(), gpu.parameters()):
        assert (pc.grad is None) == (pg.grad is None)
        if pc.grad is None:
            continue
        worst = max(worst, float((pg.grad.cpu() - pc.grad).norm()) / max(float(pc.grad.norm()), 1e-9))
    assert worst <= 2e-3, worst


def test_modelnet40_model_matches_cpu_restatement():
    """GraphConvPoolModelV1 (BASELINE configs[1], the anchor-weight family on a7-a10) on the CUDA ops vs the same network
    over the CPU oracle layers + stock-indexing conv ops: hierarchy indices bit-exact, covariance features 1e-4, loss
    1e-4, logits 1e-3, weight gradients 2e-3 Frobenius-relative."""
    from pointcloudsegmentation_b200.modelnet import GraphConvPoolModelV1
    from test_network_shapes import _modelnet_cpu, _modelnet_pair_inputs

    cpu = _modelnet_cpu()
    x, b, n, lab = _modelnet_pair_inputs()
    gpu = GraphConvPoolModelV1().cuda().eval()
    gpu.load_state_dict(cpu.state_dict())
    pc, pg = cpu.prepare(x, b, n), gpu.prepare(x.cuda(), b.cuda(), n)
    for lc, lg in zip(pc[3] + pc[4] + pc[5], pg[3] + pg[4] + pg[5]):  # vlens, vbegs, vcens
        assert torch.equal(lg.cpu(), lc)
    for gc, gg in zip(pc[6], pg[6]):  # one radius graph per level
        for a, c in zip(gg, gc):
            assert torch.equal(a.cpu(), c)
    assert torch.allclose(pg[2].cpu(), pc[2], atol=1e-4)  # permuted covariance features
    # the network comparison is fed the SAME covariance features (the fp32 kernel and the float64 restatement differ by
    # up to 1e-4 on unit-norm entries, tests/test_gpu_point_util.py; near-degenerate 2-3 point neighbourhoods amplify that
    # into the first layer's gradients)
    pg = list(pg)
    pg[2] = pc[2].cuda()
    lc, logits_c = cpu(labels=lab, prepared=pc)
    lg, logits_g = gpu(labels=lab.cuda(), prepared=pg)
    assert abs(float(lg.detach()) - float(lc.detach())) <= 1e-4 * max(1.0, abs(float(lc.detach())))
    assert torch.allclose(logits_g.cpu(), logits_c, rtol=1e-3, atol=1e-3)
    lc.backward()
    lg.backward()
    # A pre-activation within ~1e-8 of zero flips its ReLU mask between two fp32 evaluation orders (seen on this input:
    # +3.7e-9 on the CPU, -2.0e-9 on the GPU, one element of stage2.gfcs[0]) and moves that layer's weight and bias
    # gradient by ~2 % and the layers upstream of it by up to 4e-3, everything else agrees to 1e-3: allow one such layer
    # (weight + bias) 5e-2, hold the rest to 5e-3.
    errs = []
    for qc, qg in zip(cpu.parameters(), gpu.parameters()):
        assert (qc.grad is None) == (qg.grad is None)
        if qc.grad is None:
            continue
        errs.append(float((qg.grad.cpu() - qc.grad).norm()) / max(float(qc.grad.norm()), 1e-9))
    errs.sort()
    assert errs[-3] <= 5e-3 and errs[-1] <= 5e-2, errs[-4:]  # upstream of the flipped layer: <= 4e-3 observed


def test_pgnet_tower_matches_cpu_restatement():
    """PGNetV3 (model_pgnet's edge-conditioned-diffusion tower) on the CUDA ops vs the same network over the CPU oracle
    layers + stock-indexing ops: same weights, a 2-block batch; labels bit-exact, loss 1e-4, logits 1e-3, weight
    gradients 5e-3 Frobenius-relative (one ReLU-mask flip allowed 5e-2, as in the ModelNet40 test)."""
    from pointcloudsegmentation_b200.pgnet import PGNetV3
    from test_network_shapes import _pgnet_cpu

    cpu = _pgnet_cpu()
    d = s3dis_batch(9, 2, 700, 900)
    a = [torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels", "block_ids")]
    gpu = PGNetV3().cuda().eval()
    gpu.load_state_dict(cpu.state_dict())
    lc, logits_c, lab_c = cpu(*a, 2)
    lg, logits_g, lab_g = gpu(*(t.cuda() for t in a), 2)
    assert torch.equal(lab_g.cpu(), lab_c)
    assert abs(float(lg.detach()) - float(lc.detach())) <= 1e-4 * max(1.0, abs(float(lc.detach())))
    assert torch.allclose(logits_g.cpu(), logits_c, rtol=1e-3, atol=1e-3)
    lc.backward()
    lg.backward()
    errs = []
    for qc, qg in zip(cpu.parameters(), gpu.parameters()):
        assert (qc.grad is None) == (qg.grad is None)
        if qc.grad is None:
            continue
        errs.append(float((qg.grad.cpu() - qc.grad).norm()) / max(float(qc.grad.norm()), 1e-9))
    errs.sort()
    assert errs[-3] <= 5e-3 and errs[-1] <= 5e-2, errs[-4:]

"""CPU: the PyTorch restatement of pointnet_13_dilated_embed has the reference's column counts at every
stage boundary (model_pointnet.py:930-1037, counted from the concat lines): stage0 131, stage1 518,
stage2 697, global-MLP input 3 + 633 (the features BEFORE feats12 is appended, :1025-1028), lf0 1474."""
from __future__ import annotations

import torch

from oracle import cpu_layers
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed
from pointcloudsegmentation_b200.synthetic import s3dis_batch


def test_stage_widths_match_reference():
    torch.manual_seed(0)
    net = Pointnet13DilatedEmbed(L=cpu_layers.layers(), split=False).eval()
    d = s3dis_batch(3, 1, 600, 700)
    x, f, l = (torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels"))
    assert f.shape[1] == 3  # rgb, as train_graph_pool_new.py feeds it
    with torch.no_grad():
        xs, dxs, pf, lab, vl, vb, vc, graphs = net.prepare(x, f, l)
        lf2, stage1, stage0 = net.features(xs, dxs, pf, vl, vb, vc, graphs)
        lf0 = net.upsample(lf2, stage1, stage0, vl, vb, vc)
    assert stage0.shape[1] == 3 + 4 * 32 == 131
    assert stage1.shape[1] == (3 + 131 + 32) + 64 + 6 * 48 == 518
    assert lf2.shape[1] == ((3 + 518 + 48) + 64 + 64) + 128 == 697 + 128
    assert net.global_mlp.fcs[0].in_features == 3 + 633
    assert lf0.shape == (x.shape[0], 1474)


def _semantic3d_inputs(n_tile=150_000, lo=30.0, hi=36.0):
    """A 6 m x 6 m crop of a synthetic tile, voxel-sorted at the centre stride, with the dense r=0.25 neighbourhoods
    of the 0.15 m grid centres (what bench.py --dataset semantic3d prepares on the GPU)."""
    import numpy as np

    import bench
    from oracle import point_util_oracle as O

    t = bench.semantic3d_tile(0, n_tile)
    m = (t["xyzs"][:, 0] >= lo) & (t["xyzs"][:, 0] < hi) & (t["xyzs"][:, 1] >= lo) & (t["xyzs"][:, 1] < hi)
    x = t["xyzs"][m] - t["xyzs"][m].min(axis=0, keepdims=True)
    f, l = t["feats"][m], t["labels"][m]
    v = np.floor((x - x.min(0)) / np.float32(0.15)).astype(np.int64)
    o2p = np.argsort((v[:, 0] << 32) + (v[:, 1] << 16) + v[:, 2], kind="stable")
    xs, fs, ls = x[o2p], f[o2p], l[o2p]
    idxs = O.grid_downsample(xs, 0.15)
    csr = O.cross_search(xs, xs[idxs], 0.25 * 0.25)
    return [np.ascontiguousarray(a) for a in (xs, fs, ls, idxs, *csr)]


def test_semantic3d_stage_widths_match_reference():
    """pointnet_13_dilate_embed_semantic3d (model_pointnet_semantic3d.py:325-437) behind dense_feats (:307-323), counted
    from the concat lines: dense_feats 48 + 4, stage0 180, stage1 616, stage2 956, global-MLP input 3 + 828, lf0 1880."""
    from pointcloudsegmentation_b200.network import Semantic3DDense

    torch.manual_seed(0)
    net = Semantic3DDense(L=cpu_layers.layers(), split=False).eval()
    a = [torch.from_numpy(v) for v in _semantic3d_inputs()]
    with torch.no_grad():
        cx, cf, cl = net.dense_feats(*a)
        assert cf.shape == (a[3].shape[0], 48 + 4) and cx.shape[1] == 3
        xs, dxs, pf, lab, vl, vb, vc, graphs = net.prepare_dense(*a)
        lf2, stage1, stage0 = net.features(xs, dxs, pf, vl, vb, vc, graphs)
        lf0 = net.upsample(lf2, stage1, stage0, vl, vb, vc)
        loss, logits, tgt, unknown = net(*a)
    assert stage0.shape[1] == 52 + 4 * 32 == 180
    assert stage1.shape[1] == (52 + 180 + 32) + 64 + 6 * 48 == 616
    assert lf2.shape[1] == ((52 + 616 + 64) + 96 + 128) + 128 == 956 + 128
    assert net.global_mlp.fcs[0].in_features == 3 + 828
    assert lf0.shape == (a[3].shape[0], 1880)
    assert logits.shape == (a[3].shape[0], 8) and bool(torch.isfinite(loss))
    assert int(tgt.min()) >= 0 and int(tgt.max()) <= 7 and bool(unknown.any())


def _modelnet_pair_inputs(clouds=3, pn=256):
    from pointcloudsegmentation_b200.synthetic import modelnet_batch

    d = modelnet_batch(5, clouds, pn)
    return torch.from_numpy(d["xyzs"]), torch.from_numpy(d["block_ids"]), clouds, torch.tensor([3, 17, 39][:clouds])


def _modelnet_cpu():
    import sys

    from oracle import torch_ops
    from pointcloudsegmentation_b200.modelnet import GraphConvPoolModelV1

    torch.manual_seed(0)
    return GraphConvPoolModelV1(L=cpu_layers.layers(), PU=torch_ops.point_util, ops=sys.modules["oracle.torch_ops"]).eval()


def test_modelnet40_widths_match_reference():
    """graph_conv_pool_model_v1 (model.py:1193-1228) with 9 covariance features: per-stage local widths
    feats + 8 + sum(fc_dims) = 73 / 328 / 648, global feature 32 + 128 + 256 + 73 + 328 + 648, 26 anchors."""
    net = _modelnet_cpu()
    x, b, n, lab = _modelnet_pair_inputs()
    assert net.pmiu.shape == (3, 26)
    assert torch.allclose(net.pmiu.norm(dim=0), torch.ones(26), atol=1e-6)
    assert (net.stage0.out_dim, net.stage1.out_dim, net.stage2.out_dim) == (73, 328, 648)
    with torch.no_grad():
        prep = net.prepare(x, b, n)
        g = net.features(prep)
        loss, logits = net(x, b, n, lab)
    assert g.shape == (n, 32 + 128 + 256 + 73 + 328 + 648)
    assert logits.shape == (n, 40) and bool(torch.isfinite(loss))


def _pgnet_cpu():
    import sys

    from oracle import torch_ops  # noqa: F401
    from pointcloudsegmentation_b200.pgnet import PGNetV3

    torch.manual_seed(0)
    return PGNetV3(L=cpu_layers.layers(), ops=sys.modules["oracle.torch_ops"]).eval()


def test_pgnet_widths_match_reference():
    """pgnet_model_v3 (model_pgnet.py:158-205) with rgb features: stage locals 3+16+32 = 51, (64+51)+16+288 = 419,
    (256+419)+16+288 = 979; upf0 = 2*(512+979) + 256+419 + 64+51, local = 64+51."""
    net = _pgnet_cpu()
    assert (net.stage0.out_dim, net.stage1.out_dim, net.stage2.out_dim) == (51, 419, 979)
    d = s3dis_batch(3, 2, 500, 700)
    a = [torch.from_numpy(d[k]) for k in ("xyzs", "feats", "labels", "block_ids")]
    with torch.no_grad():
        prep = net.prepare(*a, 2)
        g, l = net.features(prep)
        loss, logits, lab = net(prepared=prep)
    assert g.shape == (a[0].shape[0], 2 * (512 + 979) + 256 + 419 + 64 + 51) and l.shape[1] == 64 + 51
    assert logits.shape == (a[0].shape[0], 13) and bool(torch.isfinite(loss))
    # one block alone = its slice of the batch (block-diagonal): the per-block global max / mean included
    m = a[3] == 1
    with torch.no_grad():
        _, logits1, lab1 = net(a[0][m], a[1][m], a[2][m])
    sel = torch.from_numpy(d["block_ids"]).bincount()
    assert logits1.shape[0] == int(sel[1])

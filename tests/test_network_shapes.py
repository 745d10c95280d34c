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

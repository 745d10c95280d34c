"""model_pgnet: the edge-conditioned-diffusion tower (north_star names it first among the callers of the op layer).

Restates pgnet_model_v3 (model_pgnet.py:158-205) = three ecd_stage (model_pgnet.py:65-104: search_neighborhood ->
neighbor_scatter(xyz, use_diff) / radius -> ecd_xyz -> [fc -> ecd_feats]* -> per-point MLP on [features, dxyz]) with
max / average pooling between the levels and graph_unpool back, fed by points_pooling_two_layers(0.15, 0.45, 3.0) as
conv_analysis.py:185-187 builds it, then classifier_v3 (model.py:762-799) and the weighted cross entropy of the
S3DIS training scripts.  Every custom op runs in this repo's CUDA kernels through the layer namespace `L` and the op
namespace `ops` (a1 scatter copy / diff, a3 sum gather, a4, a5, a18); the edge MLPs are stock PyTorch on materialised
[csum, C] tensors — this tower is NOT fused.

A batch of blocks is one block-diagonal problem (block_ids); the stage-2 global max / mean over a block
(tf.reduce_max / reduce_mean over axis 0, model_pgnet.py:179-181) is a segment reduction over the block's point range.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import graph_conv_layers as GC
from .graph_conv_layers import EcdFeats, _DenseFC, _mean_gather
from .network import S3DIS_TRAIN_WEIGHTS, default_layers


class EcdXyz(nn.Module):
    """ecd_xyz (model_pgnet.py:35-62): dense-concat phi-MLP on the edge coordinates, a tanh gate of the same width
    applied to the phi features themselves, g-MLP, (1+eps)/(len+eps)-weighted sum over the neighbourhood, fc + relu."""

    def __init__(self, phi_dims, g_dims, out_dim, ops=None):
        super().__init__()
        self.ops = ops
        self.phi = _DenseFC(3, phi_dims)
        self.fc_ew = nn.Linear(self.phi.out_dim, self.phi.out_dim)
        self.g = _DenseFC(self.phi.out_dim, g_dims)
        self.fc_out = nn.Linear(self.g.out_dim, out_dim)

    def forward(self, edge_coord, nlens, nbegs, ncens):
        ops = self.ops or GC._default_ops()
        p = self.phi(edge_coord)
        ef = self.g(torch.tanh(self.fc_ew(p)) * p)
        return F.relu(self.fc_out(_mean_gather(ef, ncens, nlens, nbegs, ops, eps=1e-3)))


class EcdStage(nn.Module):
    """ecd_stage (model_pgnet.py:65-104) -> (fc_final [pn, final_dim], cfeats [pn, feats_in + xyz_dim + sum(feats_dims)])."""

    def __init__(self, feats_in, xyz_dim, feats_dims, embed_dims, final_dim, ops=None):
        super().__init__()
        self.ops = ops
        self.xyz_gc = EcdXyz([xyz_dim // 2] * 2, [xyz_dim // 2] * 2, xyz_dim, ops)
        cdim = feats_in + xyz_dim
        self.fcs, self.gcs = nn.ModuleList(), nn.ModuleList()
        for fd in feats_dims:
            self.fcs.append(nn.Linear(cdim, fd))
            self.gcs.append(EcdFeats(fd, [fd // 2] * 2, [fd // 2] * 2, fd, ops))
            cdim += fd
        self.out_dim = cdim
        self.gfc = _DenseFC(cdim + 3, embed_dims)
        self.final = nn.Linear(self.gfc.out_dim, final_dim)

    def forward(self, xyzs, dxyz, feats, radius, voxel_size, graph):
        ops = self.ops or GC._default_ops()
        nidxs, nlens, nbegs, ncens = graph
        sxyzs = ops.neighbor_scatter(xyzs, nidxs, nlens, nbegs, use_diff=True) / radius
        cfeats = torch.cat([self.xyz_gc(sxyzs, nlens, nbegs, ncens), feats], dim=1)
        for fc, gc in zip(self.fcs, self.gcs):
            cfeats = torch.cat([cfeats, gc(sxyzs, F.relu(fc(cfeats)), nidxs, nlens, nbegs, ncens)], dim=1)
        return self.final(self.gfc(torch.cat([cfeats, dxyz / voxel_size], dim=1))), cfeats


class PGNetV3(nn.Module):
    """pgnet_model_v3 (model_pgnet.py:158-205) + classifier_v3 + weighted CE, for a block-diagonal batch."""

    VOXELS = (0.15, 0.45)
    BLOCK_SIZE = 3.0
    RADII = (0.15, 0.3, 0.9)  # conv_analysis.py:187

    def __init__(self, num_classes=13, feats_dim=3, L=None, ops=None, class_weights=None, dropout_keep=0.7):
        super().__init__()
        self.L = L or default_layers()
        nine = [32] * 9
        self.stage0 = EcdStage(feats_dim, 16, [16, 16], [8, 8, 8], 64, ops)
        self.stage1 = EcdStage(64 + self.stage0.out_dim, 16, nine, [32, 32, 32], 256, ops)
        self.stage2 = EcdStage(256 + self.stage1.out_dim, 16, nine, [32, 32, 32], 512, ops)
        g2 = 512 + self.stage2.out_dim
        self.local_dim = 64 + self.stage0.out_dim
        self.global_dim = 2 * g2 + 256 + self.stage1.out_dim + self.local_dim
        self.class_mlp1 = nn.Linear(self.global_dim, 512)
        self.class_mlp2 = nn.Linear(512 + self.local_dim, 256)
        self.class_mlp3 = nn.Linear(256 + self.local_dim, num_classes)
        self.dropout_keep = dropout_keep
        w = torch.tensor(class_weights if class_weights is not None else S3DIS_TRAIN_WEIGHTS[:num_classes], dtype=torch.float32)
        self.register_buffer("class_weights", w)

    def prepare(self, xyzs, feats, labels, block_ids=None, num_blocks=1):
        L = self.L
        with torch.no_grad():
            if block_ids is None:
                xs, dxs, f, lab, vl, vb, vc = L.points_pooling_two_layers(xyzs, feats, labels, *self.VOXELS, self.BLOCK_SIZE)
                n2 = xs[2].shape[0]
                bb2 = torch.tensor([0, n2], dtype=torch.int32, device=xs[2].device)
                bbegs = [(None, None)] * 3
            else:
                xs, dxs, f, lab, vl, vb, vc, bids = L.points_pooling_two_layers(xyzs, feats, labels, *self.VOXELS,
                                                                               self.BLOCK_SIZE, block_ids=block_ids)
                bbegs = []
                for b in bids:
                    bb = L.block_begs_from_ids(b, num_blocks)
                    bbegs.append((bb, int((bb[1:] - bb[:-1]).max())))
                bb2 = bbegs[2][0]
            graphs = [L.search_neighborhood(xs[k], self.RADII[k], block_begs=bbegs[k][0], max_block_pn=bbegs[k][1]) for k in range(3)]
        return xs, dxs, f, lab, vl, vb, vc, graphs, bb2

    def features(self, prepared):
        xs, dxs, f, lab, vl, vb, vc, graphs, bb2 = prepared
        L = self.L
        fc0, lf0 = self.stage0(xs[0], dxs[0], f, self.RADII[0], self.VOXELS[0], graphs[0])
        i0 = torch.cat([L.graph_pool(fc0, vl[0], vb[0]), L.graph_avg_pool(lf0, vl[0], vb[0], vc[0])], dim=1)
        fc1, lf1 = self.stage1(xs[1], dxs[1], i0, self.RADII[1], self.VOXELS[1], graphs[1])
        i1 = torch.cat([L.graph_pool(fc1, vl[1], vb[1]), L.graph_avg_pool(lf1, vl[1], vb[1], vc[1])], dim=1)
        fc2, lf2 = self.stage2(xs[2], xs[2], i1, self.RADII[2], self.BLOCK_SIZE, graphs[2])
        # per-block global max of fc2 and mean of lf2, tiled back over the block's level-2 points
        blens, bbegs = (bb2[1:] - bb2[:-1]).contiguous(), bb2[:-1].contiguous()
        bcens = torch.repeat_interleave(torch.arange(blens.shape[0], device=bb2.device, dtype=torch.int32), blens.long())
        i2 = torch.cat([L.graph_pool(fc2.contiguous(), blens, bbegs), L.graph_avg_pool(lf2.contiguous(), blens, bbegs, bcens)], dim=1)
        upf2 = torch.cat([L.graph_unpool(i2, blens, bbegs, bcens), fc2, lf2], dim=1)
        upf1 = torch.cat([L.graph_unpool(upf2, vl[1], vb[1], vc[1]), fc1, lf1], dim=1)
        upf0 = torch.cat([L.graph_unpool(upf1, vl[0], vb[0], vc[0]), fc0, lf0], dim=1)
        return upf0, torch.cat([fc0, lf0], dim=1)

    def classify(self, gfeats, lfeats):
        """classifier_v3 (model.py:762-799)."""
        p = 1.0 - self.dropout_keep
        x = F.dropout(torch.cat([F.relu(self.class_mlp1(gfeats)), lfeats], dim=1), p, self.training)
        x = F.dropout(torch.cat([F.relu(self.class_mlp2(x)), lfeats], dim=1), p, self.training)
        return self.class_mlp3(x)

    def forward(self, xyzs=None, feats=None, labels=None, block_ids=None, num_blocks=1, prepared=None):
        if prepared is None:
            prepared = self.prepare(xyzs, feats, labels, block_ids, num_blocks)
        lab = prepared[3]
        logits = self.classify(*self.features(prepared))
        w = self.class_weights[lab]
        ce = F.cross_entropy(logits, lab, reduction="none")
        loss = (ce * w).sum() / torch.count_nonzero(w).clamp(min=1)
        return loss, logits, lab

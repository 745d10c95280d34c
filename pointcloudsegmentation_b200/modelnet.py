"""ModelNet40 classification (BASELINE configs[1]: train_modelnet40, 1024 points per cloud, batch 32): the anchor-weight
graph-convolution family on the location-weight ops (a7-a10).

Restates tower_loss of train_modelnet40.py:41-58 = graph_conv_pool_model_v1 (model.py:1193-1228: three
graph_pmiu_nosum_all_conv_pool_stage stages, model.py:1062-1088, blocks model.py:383-407) -> model_classifier_v1
(model.py:1231-1262) -> softmax cross entropy, on hierarchies built on the GPU (normalize_model_hierarchy,
aug_util.py:467-511: radii 0.05 / 0.1 / 0.3, voxels 0.1 / 0.3, covariance features) by points_pooling_two_layers,
the batched radius searches and computeCovarsGPU.

The reference feeds ONE cloud per tower; here a batch of clouds is one block-diagonal problem (block_ids), and the
per-cloud global max (tf.reduce_max(.., axis=0), model.py:1217-1223) is a segment max over the clouds' point ranges.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import graph_conv_layers as GC


def neighbor_anchors_v2() -> torch.Tensor:
    """train_modelnet40.py:215-227 -> pmiu [3, 26]: 3 x 8 directions at elevations -45, 0, 45 degrees + the two poles."""
    interval = 2 * np.pi / 8
    pmiu = []
    for va in np.arange(-np.pi / 2 + interval, np.pi / 2, interval):
        for ha in np.arange(0, 2 * np.pi, interval):
            pmiu.append([np.cos(va) * np.cos(ha), np.cos(va) * np.sin(ha), np.sin(va)])
    pmiu += [[0, 0, 1.0], [0, 0, -1.0]]
    return torch.tensor(np.asarray(pmiu, dtype=np.float32).T.copy())


def _xavier(*shape):
    p = nn.Parameter(torch.empty(*shape))
    nn.init.xavier_uniform_(p.view(shape[0], -1))
    return p


class ConvPoolStage(nn.Module):
    """graph_pmiu_nosum_all_conv_pool_stage (model.py:1062-1088): anchor-weighted xyz convolution (computes the location
    weights lw, lw_sum of the stage), a chain of anchor-weighted feature convolutions with dense concatenation, then a
    per-point MLP on [features, dxyz].  -> (fc_final [pn, final_dim], cfeats [pn, feats_dim + gxyz_dim + sum(fc_dims)])."""

    def __init__(self, feats_dim, gxyz_dim, gc_dims, fc_dims, gfc_dims, final_dim, m, ops=None):
        super().__init__()
        self.ops = ops
        self.pw_xyz = _xavier(3, m, gxyz_dim)
        self.b_xyz = nn.Parameter(torch.zeros(m * gxyz_dim))
        self.xyz_fc = nn.Linear(m * gxyz_dim, gxyz_dim)
        cdim = feats_dim + gxyz_dim
        self.pws, self.biases, self.fcs = nn.ParameterList(), nn.ParameterList(), nn.ModuleList()
        for gd, fd in zip(gc_dims, fc_dims):
            self.pws.append(_xavier(cdim, m, gd))
            self.biases.append(nn.Parameter(torch.zeros(m * gd)))
            self.fcs.append(nn.Linear(m * gd + cdim, fd))
            cdim += fd
        self.out_dim = cdim
        dims = [cdim + 3] + list(gfc_dims)
        self.gfcs = nn.ModuleList([nn.Linear(a, b) for a, b in zip(dims[:-1], dims[1:])])
        self.final = nn.Linear(dims[-1], final_dim)

    def forward(self, cxyz, dxyz, feats, cidxs, nidxs, nlens, nbegs, pmiu):
        kw = dict(ops=self.ops) if self.ops is not None else {}
        g, lw, lw_sum = GC.graph_conv_xyz_impl(cxyz, cidxs, nidxs, nlens, nbegs, self.pw_xyz, compute_lw=True, pmiu=pmiu,
                                               no_sum=True, **kw)
        g = F.relu(self.xyz_fc(F.relu(g + self.b_xyz)))
        cfeats = torch.cat([g, feats], dim=1)
        for pw, b, fc in zip(self.pws, self.biases, self.fcs):  # graph_diff_conv_pool_block, use_diff=False
            gc = F.relu(GC.graph_conv_feats_impl(cfeats, cidxs, nidxs, nlens, nbegs, pw, lw, lw_sum, no_sum=True, **kw) + b)
            cfeats = torch.cat([F.relu(fc(torch.cat([gc, cfeats], dim=1))), cfeats], dim=1)
        x = torch.cat([cfeats, dxyz], dim=1)
        for l in self.gfcs:
            x = F.relu(l(x))
        return self.final(x), cfeats


class GraphConvPoolModelV1(nn.Module):
    """graph_conv_pool_model_v1 (model.py:1193-1228) + model_classifier_v1 (model.py:1231-1262) + the loss of
    train_modelnet40.py:52-54, for a batch of clouds."""

    RADII = (0.05, 0.1, 0.3)   # aug_util.py:467 nr1, nr2, nr3
    VOXELS = (0.1, 0.3)        # vc1, vc2

    def __init__(self, num_classes=40, feats_dim=9, L=None, PU=None, ops=None, dropout_keep=0.7):
        super().__init__()
        if L is None:
            from . import graph_layers as L
        self.L, self.PU = L, PU
        m = 26
        self.register_buffer("pmiu", neighbor_anchors_v2())
        self.stage0 = ConvPoolStage(feats_dim, 8, [8, 16, 32], [8, 16, 32], [32, 32, 32], 32, m, ops)
        self.stage1 = ConvPoolStage(32, 8, [32, 32, 32, 64, 64, 64], [32, 32, 32, 64, 64, 64], [128, 128, 128], 128, m, ops)
        self.stage2 = ConvPoolStage(128, 8, [128, 128, 256], [128, 128, 256], [256, 256], 256, m, ops)
        gdim = 32 + 128 + 256 + self.stage0.out_dim + self.stage1.out_dim + self.stage2.out_dim
        self.class_fc1 = nn.Linear(gdim, 512)
        self.class_fc2 = nn.Linear(512 + gdim, 256)
        self.class_fc3 = nn.Linear(256 + gdim, num_classes)
        self.dropout_keep = dropout_keep

    def prepare(self, xyzs, block_ids, num_blocks):
        """normalize_model_hierarchy (aug_util.py:467-511) for a batch: covariance features of the r = 0.05
        neighbourhoods, two voxel levels (dxyz rescaled to the voxel), one radius graph per level."""
        L = self.L
        with torch.no_grad():
            xyzs = xyzs.contiguous()
            bb0 = L.block_begs_from_ids(block_ids, num_blocks)
            mb0 = int((bb0[1:] - bb0[:-1]).max())
            n0 = L.search_neighborhood(xyzs, self.RADII[0], block_begs=bb0, max_block_pn=mb0)
            if self.PU is not None:
                covars = self.PU.computeCovarsGPU(xyzs, n0[0], n0[1], n0[2])
            else:
                from . import point_util

                covars = point_util.computeCovarsGPU(xyzs, n0[0], n0[1], n0[2])
            labels = torch.zeros(xyzs.shape[0], dtype=torch.int64, device=xyzs.device)
            # clouds live in the unit sphere; compute_voxel_index puts the grid's corner at (-bs/2, -bs/2, 0): pool
            # the cloud lifted by one unit in z with block size 2, then take the lift out of the coordinates again
            # (everything but stage 2's absolute-coordinate input is translation invariant)
            lift = torch.tensor([0.0, 0.0, 1.0], dtype=xyzs.dtype, device=xyzs.device)
            xs, dxs, f, _, vl, vb, vc, bids = L.points_pooling_two_layers(xyzs + lift, covars, labels, *self.VOXELS, 2.0,
                                                                          block_ids=block_ids)
            xs = [(x - lift).contiguous() for x in xs]
            dxs = [dxs[0] / self.VOXELS[0], dxs[1] / self.VOXELS[1]]
            graphs, bbs = [], []
            for lvl in range(3):
                bb = L.block_begs_from_ids(bids[lvl], num_blocks)
                mb = int((bb[1:] - bb[:-1]).max())
                bbs.append(bb)
                graphs.append(L.search_neighborhood(xs[lvl], self.RADII[lvl], block_begs=bb, max_block_pn=mb))
        return xs, dxs, f, vl, vb, vc, graphs, bbs

    def features(self, prepared):
        xs, dxs, f, vl, vb, vc, graphs, bbs = prepared
        L = self.L
        g = lambda k: (graphs[k][3], graphs[k][0], graphs[k][1], graphs[k][2])  # noqa: E731  cidxs, nidxs, lens, begs
        fc0, lf0 = self.stage0(xs[0], dxs[0], f, *g(0), self.pmiu)
        fc1, lf1 = self.stage1(xs[1], dxs[1], L.graph_pool(fc0, vl[0], vb[0]), *g(1), self.pmiu)
        fc2, lf2 = self.stage2(xs[2], xs[2], L.graph_pool(fc1, vl[1], vb[1]), *g(2), self.pmiu)
        # per-cloud global max, one segment max per level over [fc | lf] (a cloud is a contiguous point range)
        outs = [self._cloud_max(torch.cat([a, b], dim=1), bbs[lvl]) for lvl, (a, b) in enumerate(((fc0, lf0), (fc1, lf1), (fc2, lf2)))]
        w = (fc0.shape[1], fc1.shape[1], fc2.shape[1])
        return torch.cat([o[:, :k] for o, k in zip(outs, w)] + [o[:, k:] for o, k in zip(outs, w)], dim=1)

    def _cloud_max(self, t, bb, chunk=32):
        """max over every cloud's rows [bb[c], bb[c+1]): two segment-max passes (runs of <= `chunk` rows, then the runs of a
        cloud) so that a 1024-row cloud is not one serial walk per column."""
        L = self.L
        lens = (bb[1:] - bb[:-1]).long()
        nch = (lens + chunk - 1) // chunk                                   # runs per cloud
        run_cloud = torch.repeat_interleave(torch.arange(lens.shape[0], device=t.device), nch)
        run_first = torch.cumsum(nch, 0) - nch
        k = torch.arange(run_cloud.shape[0], device=t.device) - run_first[run_cloud]   # run index inside its cloud
        rbeg = bb[:-1].long()[run_cloud] + k * chunk
        rlen = torch.minimum(torch.full_like(rbeg, chunk), bb[1:].long()[run_cloud] - rbeg)
        part = L.graph_pool(t.contiguous(), rlen.to(torch.int32), rbeg.to(torch.int32))
        return L.graph_pool(part, nch.to(torch.int32), run_first.to(torch.int32))

    def classify(self, feats):
        p = 1.0 - self.dropout_keep
        x = torch.cat([F.relu(self.class_fc1(feats)), feats], dim=1)
        x = F.dropout(x, p, self.training)
        x = torch.cat([F.relu(self.class_fc2(x)), feats], dim=1)
        x = F.dropout(x, p, self.training)
        return self.class_fc3(x)

    def forward(self, xyzs=None, block_ids=None, num_blocks=1, labels=None, prepared=None):
        """xyzs [B*pn, 3] (every cloud inside the unit sphere), block_ids [B*pn] ascending, labels [B]
        -> (loss, logits [B, classes])."""
        if prepared is None:
            prepared = self.prepare(xyzs, block_ids, num_blocks)
        logits = self.classify(self.features(prepared))
        return F.cross_entropy(logits, labels), logits

"""TEST INFRASTRUCTURE, NOT PRODUCT.  The layer namespace of pointcloudsegmentation_b200/network.py
re-implemented on CPU torch tensors over the C oracle (oracle/pcs_oracle.py): what the reference's
op layer computes, on the host cores.  Used by (1) tests/ to check the GPU network's loss and
gradients end to end, (2) bench.py's `cpu_baseline` leg and `--impl reference` arm (the reference has
no CPU implementation of these ops — every kernel is DEVICE_GPU only — so the oracle port is the CPU
path that gets timed; `kind` = "port").  Gradient wiring mirrors tf_ops/graph_conv_layer.py:11-45 and
tf_ops/graph_pooling_layer.py:11-18.  Sums are returned rounded to fp32 from the oracle's float64.
"""
from __future__ import annotations

import types

import numpy as np
import torch

from . import pcs_oracle as orc


def _n(t):
    return t.detach().cpu().numpy()


def _t(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t if dtype is None else t.to(dtype)


class _Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ifeats, nidxs, lens, bgs, use_diff):
        ctx.a = (_n(nidxs), _n(lens), _n(bgs), bool(use_diff))
        return _t(orc.neighbor_scatter(_n(ifeats), *ctx.a[:3], use_diff=ctx.a[3]))

    @staticmethod
    def backward(ctx, d):
        nidxs, lens, bgs, ud = ctx.a
        return _t(orc.neighbor_gather(_n(d), nidxs, lens, bgs, ud), torch.float32), None, None, None, None


class _SumGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, cidxs, lens, bgs):
        ctx.a = (_n(cidxs), _n(lens), _n(bgs))
        return _t(orc.segment_sum(_n(feats), ctx.a[1], ctx.a[2]), torch.float32)

    @staticmethod
    def backward(ctx, d):
        return _t(orc.row_gather(_n(d).astype(np.float32), ctx.a[0])), None, None, None


class _SumScatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, cidxs, lens, bgs):
        ctx.a = (_n(cidxs), _n(lens), _n(bgs))
        return _t(orc.row_gather(_n(feats), ctx.a[0]))

    @staticmethod
    def backward(ctx, d):
        return _t(orc.segment_sum(_n(d), ctx.a[1], ctx.a[2]), torch.float32), None, None, None


class _MaxGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, vlens, vbegs):
        lens, bgs = _n(vlens), _n(vbegs)
        g, mi = orc.max_feat_gather(_n(feats), lens, bgs)
        ctx.a = (mi, lens, bgs, feats.shape[0])
        return _t(g)

    @staticmethod
    def backward(ctx, d):
        mi, lens, bgs, pn1 = ctx.a
        return _t(orc.max_feat_scatter(_n(d), mi, bgs, lens, pn1)), None, None


def neighbor_scatter(ifeats, nidxs, lens, bgs, use_diff=False):
    return _Scatter.apply(ifeats, nidxs, lens, bgs, use_diff)


def graph_pool(feats, vlens, vbegs):
    return _MaxGather.apply(feats, vlens, vbegs)


def graph_unpool(feats, vlens, vbegs, cidxs):
    return _SumScatter.apply(feats, cidxs, vlens, vbegs)


def graph_avg_pool(feats, vlens, vbegs, vcens):
    return _SumGather.apply(feats, vcens, vlens, vbegs) / vlens.to(torch.float32).unsqueeze(1)


def graph_concat_scatter(feats, nidxs, nlens, nbegs, ncens):
    return torch.cat([graph_unpool(feats, nlens, nbegs, ncens), neighbor_scatter(feats, nidxs, nlens, nbegs, False)], 1)


def search_neighborhood(xyzs, radius, block_begs=None, max_block_pn=None):
    bb = None if block_begs is None else _n(block_begs)
    return tuple(_t(a) for a in orc.search(_n(xyzs), radius * radius, None, bb))


def search_neighborhood_range(xyzs, min_radius, max_radius, block_begs=None, max_block_pn=None):
    bb = None if block_begs is None else _n(block_begs)
    return tuple(_t(a) for a in orc.search(_n(xyzs), max_radius * max_radius, min_radius * min_radius, bb))


def block_begs_from_ids(sorted_block_ids, num_blocks):
    counts = np.bincount(_n(sorted_block_ids), minlength=num_blocks)
    return _t(np.concatenate([[0], np.cumsum(counts)]).astype(np.int32))


def _avg(pts, cens, lens, begs):
    # fp32 division of the fp32-rounded sum, as the TF graph does (graph_pooling_layer.py:124-125)
    s = orc.segment_sum(pts, lens, begs).astype(np.float32)
    return s / lens.astype(np.float32)[:, None]


def points_pooling_two_layers(xyzs, feats, labels, voxel_size1, voxel_size2, block_size=3.0, block_ids=None):
    """tf_ops/graph_pooling_layer.py:110-146 on numpy; `feats` stays differentiable through the two
    row permutations (PermutateFeature -> PermutateFeatureBackward)."""
    x = _n(xyzs)
    lab = _n(labels).astype(np.int32)[:, None]
    bids = None if block_ids is None else _n(block_ids)
    o2p1, p2o1, vl1, vb1, vc1 = orc.permutation_info(orc.voxel_index(x, voxel_size1, block_size), None, bids)
    pts1 = orc.row_gather(x, o2p1)
    lab = orc.row_gather(lab, o2p1)
    b1 = None if bids is None else bids[o2p1]
    b2 = None if bids is None else b1[vb1]
    pts2 = _avg(pts1, vc1, vl1, vb1)
    dpts1 = orc.diff_xyz(pts1, pts2, vc1)
    o2p2, p2o2, vl2, vb2, vc2 = orc.permutation_info(orc.voxel_index(pts2, voxel_size2, block_size), None, b2)
    pts2 = orc.row_gather(pts2, o2p2)
    if bids is not None:
        b2 = b2[o2p2]
    b3 = None if bids is None else b2[vb2]
    pts3 = _avg(pts2, vc2, vl2, vb2)
    dpts2 = orc.diff_xyz(pts2, pts3, vc2)
    ro2p, rp2o, vl1, vb1, vc1 = orc.repermutation_info(o2p2, vl1, vb1, x.shape[0])
    pts1 = orc.row_gather(pts1, ro2p)
    dpts1 = orc.row_gather(dpts1, ro2p)
    lab = orc.row_gather(lab, ro2p)
    if bids is not None:
        b1 = b1[ro2p]
    total = torch.from_numpy(o2p1[ro2p].astype(np.int64))  # the two permutations composed
    f = feats[total]  # torch indexing: differentiable, same values as two PermutateFeature calls
    out = ([_t(pts1), _t(pts2), _t(pts3)], [_t(dpts1), _t(dpts2)], f, _t(lab[:, 0]).to(torch.int64),
           [_t(vl1), _t(vl2)], [_t(vb1), _t(vb2)], [_t(vc1), _t(vc2)])
    if bids is None:
        return out
    return out + ([_t(b1), _t(b2), _t(b3)],)


def layers():
    return types.SimpleNamespace(
        points_pooling_two_layers=points_pooling_two_layers, search_neighborhood=search_neighborhood,
        search_neighborhood_range=search_neighborhood_range, graph_avg_pool=graph_avg_pool,
        graph_concat_scatter=graph_concat_scatter, graph_pool=graph_pool, graph_unpool=graph_unpool,
        neighbor_scatter=neighbor_scatter, block_begs_from_ids=block_begs_from_ids)

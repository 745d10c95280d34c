"""`neighbor_ops` — the namespace the reference gets from
`tf.load_op_library('build/libTFNeighborOps.so')` (tf_ops/graph_conv_layer.py:4), with the gradient
wiring of its `@ops.RegisterGradient` stubs (tf_ops/graph_conv_layer.py:11-45) expressed as
`torch.autograd.Function`s.  Same op names, same positional argument order, same outputs.

Index tensors (int32) never carry gradients; backward-only ops (`neighbor_gather`,
`*_backward`, `neighbor_max_feat_scatter`) are exported raw, as in the reference.
"""
from __future__ import annotations

import torch

from . import ops as _raw

# backward-only / integer ops: exported as-is (the reference registers no gradient for them)
neighbor_gather = _raw.neighbor_gather
neighbor_max_feat_scatter = _raw.neighbor_max_feat_scatter
location_weight_sum_backward = _raw.location_weight_sum_backward
location_weight_feat_sum_backward = _raw.location_weight_feat_sum_backward
eliminate_center = _raw.eliminate_center


class _NeighborScatter(torch.autograd.Function):
    # gradient: NeighborGather with the same use_diff (graph_conv_layer.py:11-15)
    @staticmethod
    def forward(ctx, ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff):
        ctx.csr = (nidxs, nidxs_lens, nidxs_bgs)
        ctx.use_diff = bool(use_diff)
        return _raw.neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff)

    @staticmethod
    def backward(ctx, dsfeats):
        nidxs, lens, bgs = ctx.csr
        return _raw.neighbor_gather(dsfeats.contiguous(), nidxs, lens, bgs, ctx.use_diff), None, None, None, None


class _NeighborSumFeatGather(torch.autograd.Function):
    # gradient: NeighborSumFeatScatter (graph_conv_layer.py:24-27)
    @staticmethod
    def forward(ctx, ifeats, cidxs, nidxs_lens, nidxs_bgs):
        ctx.csr = (cidxs, nidxs_lens, nidxs_bgs)
        return _raw.neighbor_sum_feat_gather(ifeats, cidxs, nidxs_lens, nidxs_bgs)

    @staticmethod
    def backward(ctx, dgfeats):
        cidxs, lens, bgs = ctx.csr
        return _raw.neighbor_sum_feat_scatter(dgfeats.contiguous(), cidxs, lens, bgs), None, None, None


class _NeighborSumFeatScatter(torch.autograd.Function):
    # gradient: NeighborSumFeatGather (graph_conv_layer.py:30-33)
    @staticmethod
    def forward(ctx, ifeats, cidxs, nidxs_lens, nidxs_bgs):
        ctx.csr = (cidxs, nidxs_lens, nidxs_bgs)
        return _raw.neighbor_sum_feat_scatter(ifeats, cidxs, nidxs_lens, nidxs_bgs)

    @staticmethod
    def backward(ctx, dsfeats):
        cidxs, lens, bgs = ctx.csr
        return _raw.neighbor_sum_feat_gather(dsfeats.contiguous(), cidxs, lens, bgs), None, None, None


class _NeighborMaxFeatGather(torch.autograd.Function):
    # gradient: NeighborMaxFeatScatter(dg, feats, max_idxs, vbegs, vlens) (graph_conv_layer.py:42-45)
    @staticmethod
    def forward(ctx, ifeats, vlens, vlens_bgs):
        gfeats, max_idxs = _raw.neighbor_max_feat_gather(ifeats, vlens, vlens_bgs)
        ctx.saved = (max_idxs, vlens, vlens_bgs, int(ifeats.shape[0]))
        ctx.mark_non_differentiable(max_idxs)
        return gfeats, max_idxs

    @staticmethod
    def backward(ctx, dgfeats, _dmax_idxs):
        max_idxs, vlens, vbegs, pn1 = ctx.saved
        return _raw.neighbor_max_feat_scatter(dgfeats.contiguous(), pn1, max_idxs, vbegs, vlens), None, None


class _LocationWeightSum(torch.autograd.Function):
    # gradient: LocationWeightSumBackward (graph_conv_layer.py:36-39)
    @staticmethod
    def forward(ctx, lw, nidxs_lens, nidxs_bgs):
        ctx.csr = (nidxs_lens, nidxs_bgs)
        ctx.save_for_backward(lw)
        return _raw.location_weight_sum(lw, nidxs_lens, nidxs_bgs)

    @staticmethod
    def backward(ctx, dlw_sum):
        (lw,) = ctx.saved_tensors
        lens, bgs = ctx.csr
        return _raw.location_weight_sum_backward(lw, dlw_sum.contiguous(), lens, bgs), None, None


class _LocationWeightFeatSum(torch.autograd.Function):
    # gradient: LocationWeightFeatSumBackward (graph_conv_layer.py:18-21)
    @staticmethod
    def forward(ctx, lw, feats, nidxs_lens, nidxs_bgs):
        ctx.csr = (nidxs_lens, nidxs_bgs)
        ctx.save_for_backward(lw, feats)
        return _raw.location_weight_feat_sum(lw, feats, nidxs_lens, nidxs_bgs)

    @staticmethod
    def backward(ctx, dtfeats_sum):
        lw, feats = ctx.saved_tensors
        lens, bgs = ctx.csr
        dlw, dfeats = _raw.location_weight_feat_sum_backward(lw, feats, dtfeats_sum.contiguous(), lens, bgs)
        return dlw, dfeats, None, None


def neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False):
    return _NeighborScatter.apply(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff)


def neighbor_sum_feat_gather(ifeats, cidxs, nidxs_lens, nidxs_bgs):
    return _NeighborSumFeatGather.apply(ifeats, cidxs, nidxs_lens, nidxs_bgs)


def neighbor_sum_feat_scatter(ifeats, cidxs, nidxs_lens, nidxs_bgs):
    return _NeighborSumFeatScatter.apply(ifeats, cidxs, nidxs_lens, nidxs_bgs)


def neighbor_max_feat_gather(ifeats, vlens, vlens_bgs):
    return _NeighborMaxFeatGather.apply(ifeats, vlens, vlens_bgs)


def location_weight_sum(lw, nidxs_lens, nidxs_bgs):
    return _LocationWeightSum.apply(lw, nidxs_lens, nidxs_bgs)


def location_weight_feat_sum(lw, feats, nidxs_lens, nidxs_bgs):
    return _LocationWeightFeatSum.apply(lw, feats, nidxs_lens, nidxs_bgs)

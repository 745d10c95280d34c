"""TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT.  The layer namespace of network.Pointnet13DilatedEmbed backed by
the reference's OWN CUDA kernels (oracle/_ref/libref.so = /root/reference/tf_ops/**/*.cu compiled unmodified for
sm_100a, see oracle/Makefile): the "kernel to beat" of SURVEY.md §2.3 as a whole training step on the same B200.

Gradient wiring as tf_ops/graph_conv_layer.py:11-45 and tf_ops/graph_pooling_layer.py:11-18 register it; composites as
tf_ops/graph_pooling_layer.py:39-146.  One block per call, like the reference feeds one block per tower
(train_graph_pool.py:95-108): the reference kernels know nothing of block-diagonal batches.  Used only by bench.py's
`reference_cuda` block and by tests; nothing under pointcloudsegmentation_b200/ imports it.
"""
from __future__ import annotations

import types

import torch

from . import libref as R


class _Scatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ifeats, nidxs, lens, bgs, use_diff):
        ctx.save_for_backward(nidxs, lens, bgs)
        ctx.use_diff = use_diff
        return R.neighbor_scatter(ifeats.contiguous(), nidxs, lens, bgs, use_diff)

    @staticmethod
    def backward(ctx, d):
        nidxs, lens, bgs = ctx.saved_tensors
        return R.neighbor_gather(d.contiguous(), nidxs, lens, bgs, ctx.use_diff), None, None, None, None


class _SumGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, cidxs, lens, bgs):
        ctx.save_for_backward(cidxs)
        return R.sum_feat_gather(feats.contiguous(), lens, bgs)

    @staticmethod
    def backward(ctx, d):
        (cidxs,) = ctx.saved_tensors
        return R.sum_feat_scatter(d.contiguous(), cidxs), None, None, None


class _SumScatter(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, cidxs, lens, bgs):
        ctx.save_for_backward(lens, bgs)
        return R.sum_feat_scatter(feats.contiguous(), cidxs)

    @staticmethod
    def backward(ctx, d):
        lens, bgs = ctx.saved_tensors
        return R.sum_feat_gather(d.contiguous(), lens, bgs), None, None, None


class _MaxGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, vlens, vbegs):
        g, mi = R.max_feat_gather(feats.contiguous(), vlens, vbegs)
        ctx.save_for_backward(mi, vlens, vbegs)
        ctx.pn1 = feats.shape[0]
        return g

    @staticmethod
    def backward(ctx, d):
        mi, vlens, vbegs = ctx.saved_tensors
        return R.max_feat_scatter(d.contiguous(), mi, vbegs, vlens, ctx.pn1), None, None


class _Permutate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, o2p, p2o):
        ctx.save_for_backward(p2o)
        return R.permutate_feature(feats.contiguous(), o2p)

    @staticmethod
    def backward(ctx, d):
        (p2o,) = ctx.saved_tensors
        return R.permutate_feature(d.contiguous(), p2o), None, None


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
    assert block_begs is None, "the reference searches one block per call"
    return R.search(xyzs, radius * radius)  # the attr is squared in Python and cast to float by the op (.cc:35)


def search_neighborhood_range(xyzs, min_radius, max_radius, block_begs=None, max_block_pn=None):
    assert block_begs is None, "the reference searches one block per call"
    return R.search(xyzs, max_radius * max_radius, min_radius * min_radius)


def _voxel_index(pts, voxel_len, block_size):
    # lower corner formed in fp32 as the op does (PoolingIndex/ComputeVoxelIdx.cc:66-67)
    import numpy as np

    f = np.float32
    eps = f(1e-3)
    mn_xy = -f(block_size) / f(2.0) - eps
    return R.voxel_index(pts, float(f(voxel_len)), float(mn_xy), float(mn_xy), float(f(0.0) - eps))


def points_pooling_two_layers(xyzs, feats, labels, voxel_size1, voxel_size2, block_size=3.0, block_ids=None):
    """tf_ops/graph_pooling_layer.py:110-146 on the reference kernels, one block."""
    assert block_ids is None, "the reference pools one block per call"
    lab = labels.to(torch.int32).unsqueeze(1).to(torch.float32)
    o2p1, p2o1, vl1, vb1, vc1 = R.permutation_info(_voxel_index(xyzs, voxel_size1, block_size))
    pts1 = R.permutate_feature(xyzs, o2p1)
    lab = R.permutate_feature(lab, o2p1)
    f = _Permutate.apply(feats, o2p1, p2o1)
    pts2 = R.sum_feat_gather(pts1, vl1, vb1) / vl1.to(torch.float32).unsqueeze(1)
    dpts1 = R.diff_xyz(pts1, pts2, vc1)
    o2p2, p2o2, vl2, vb2, vc2 = R.permutation_info(_voxel_index(pts2, voxel_size2, block_size))
    pts2 = R.permutate_feature(pts2, o2p2)
    pts3 = R.sum_feat_gather(pts2, vl2, vb2) / vl2.to(torch.float32).unsqueeze(1)
    dpts2 = R.diff_xyz(pts2, pts3, vc2)
    ro2p, rp2o, vl1, vb1, vc1 = R.repermutation_info(o2p2, vl1, vb1, xyzs.shape[0])
    pts1 = R.permutate_feature(pts1, ro2p)
    dpts1 = R.permutate_feature(dpts1, ro2p)
    lab = R.permutate_feature(lab, ro2p)
    f = _Permutate.apply(f, ro2p, rp2o)
    return ([pts1, pts2, pts3], [dpts1, dpts2], f, lab[:, 0].to(torch.int64), [vl1, vl2], [vb1, vb2], [vc1, vc2])


def layers():
    return types.SimpleNamespace(
        points_pooling_two_layers=points_pooling_two_layers, search_neighborhood=search_neighborhood,
        search_neighborhood_range=search_neighborhood_range, graph_avg_pool=graph_avg_pool,
        graph_concat_scatter=graph_concat_scatter, graph_pool=graph_pool, graph_unpool=graph_unpool,
        neighbor_scatter=neighbor_scatter)

"""`pooling_ops` — the namespace the reference gets from `tf.load_op_library('build/PoolingOps.so')`
(tf_ops/graph_pooling_layer.py:4) with its gradient registrations (graph_pooling_layer.py:11-18):
PermutateFeature -> PermutateFeatureBackward(d, back_idxs); ComputeDiffXyz -> no gradient at all.
TF attrs (`voxel_len=`, `squared_nn_size=`, ...) are keyword arguments here too.
"""
from __future__ import annotations

import torch

from . import ops as _raw

compute_voxel_index = _raw.compute_voxel_index
compute_permutation_info = _raw.compute_permutation_info
compute_permutation_info_with_class = _raw.compute_permutation_info_with_class
compute_repermutation_info = _raw.compute_repermutation_info
permutate_feature_backward = _raw.permutate_feature_backward
search_neighborhood_brute_force = _raw.search_neighborhood_brute_force
search_neighborhood_brute_force_range = _raw.search_neighborhood_brute_force_range
search_neighborhood_fixed_brute_force = _raw.search_neighborhood_fixed_brute_force
search_neighborhood_fixed_brute_force_range = _raw.search_neighborhood_fixed_brute_force_range


class _PermutateFeature(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, idxs, back_idxs):
        ctx.back_idxs = back_idxs
        return _raw.permutate_feature(feats, idxs, back_idxs)

    @staticmethod
    def backward(ctx, dfeats):
        return _raw.permutate_feature_backward(dfeats.contiguous(), ctx.back_idxs), None, None


def permutate_feature(feats, idxs, back_idxs):
    if feats.dtype != torch.float32 or not feats.requires_grad:
        return _raw.permutate_feature(feats, idxs, back_idxs)
    return _PermutateFeature.apply(feats, idxs, back_idxs)


def compute_diff_xyz(xyzs, cxyzs, cidxs):
    # gradient registered as [None, None, None] (graph_pooling_layer.py:16-18): the result is a constant
    return _raw.compute_diff_xyz(xyzs.detach(), cxyzs.detach(), cidxs)

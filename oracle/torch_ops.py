"""TEST INFRASTRUCTURE, NOT PRODUCT.  The four differentiable ops the L3 convolution layers call
(tf_ops/graph_conv_layer.py:74-166), written with stock PyTorch indexing on any device and dtype (autograd by torch):
the restatement the anchor-weight networks (pointcloudsegmentation_b200/modelnet.py) are checked against, next to
oracle/cpu_layers.py for the pooling / search layers.  Semantics per op: neighbor_scatter TFNeighborKernel.cu:5-59,
neighbor_sum_feat_gather TFNeighborKernelNew.cu:4-27, location_weight_sum TFNeighborKernel.cu:128-151,
location_weight_feat_sum TFNeighborKernel.cu:179-211."""
from __future__ import annotations

import torch


def _cens(lens):
    return torch.repeat_interleave(torch.arange(lens.shape[0], device=lens.device), lens.long())


def neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False):
    s = ifeats[nidxs.long()]
    return s - ifeats[_cens(nidxs_lens)] if use_diff else s


def neighbor_sum_feat_gather(ifeats, cidxs, nidxs_lens, nidxs_bgs):
    out = torch.zeros(nidxs_lens.shape[0], ifeats.shape[1], dtype=ifeats.dtype, device=ifeats.device)
    return out.index_add(0, _cens(nidxs_lens), ifeats)


def location_weight_sum(lw, nidxs_lens, nidxs_bgs):
    return neighbor_sum_feat_gather(lw, None, nidxs_lens, nidxs_bgs)


def location_weight_feat_sum(lw, feats, nidxs_lens, nidxs_bgs):
    w = (lw.unsqueeze(2) * feats).reshape(feats.shape[0], -1)
    return neighbor_sum_feat_gather(w, None, nidxs_lens, nidxs_bgs).reshape(-1, *feats.shape[1:])


class _PU:
    """computeCovarsGPU of the libPointUtil restatement, on torch CPU tensors."""

    @staticmethod
    def computeCovarsGPU(xyzs, nidxs, lens, begs):
        from . import point_util_oracle as O

        c = O.compute_covars(xyzs.numpy(), nidxs.numpy(), lens.numpy(), begs.numpy())
        return torch.from_numpy(c.astype("float32"))


point_util = _PU()

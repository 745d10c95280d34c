"""L3 graph-convolution layers of the reference, on this package's ops: the callers of a1, a3, a7-a10.

Functional mirrors (weights passed in, same names / argument order / formulas) of
    tf_ops/graph_conv_layer.py:74-166   compute_pfeats, graph_conv_xyz_feats_impl, graph_conv_xyz_impl,
                                        graph_conv_feats_impl, graph_diff_conv_feats_impl
and nn.Module mirrors (own parameters, dense-concat `fully_connected` chains as in the reference) of
    tf_ops/graph_conv_layer.py:573-627  graph_conv_edge, graph_conv_edge_xyz   (edge-conditioned weights)
    model_pgnet.py:4-32                 ecd_feats                               (gated edge diffusion)
These are the layers the reference's model.py / model_pgnet.py networks are assembled from
(graph_conv_pool_stage_edge model.py:1793-1821, ecd_stage model_pgnet.py:65-105).

`fused=True` in compute_pfeats uses LocationWeightFeatSum (a9) — the op the reference has commented out at
graph_conv_layer.py:93 — instead of materialising lw*tfeats [csum, m*ofn] and summing it (a3): same sums,
one pass over tfeats instead of three.

`ops` is the op namespace: default = this package's CUDA ops (neighbor_ops); tests pass a plain-PyTorch
namespace to check the wiring in float64.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _default_ops():
    from . import neighbor_ops

    return neighbor_ops


def compute_pfeats(lw, lw_sum, tfeats, nidxs_lens, nidxs_bgs, cidxs, no_sum=False, fused=True, ops=None):
    """lw [csum,m], lw_sum [pn,m], tfeats [csum,m,ofn] -> pfeats [pn,ofn] (or [pn,m*ofn] if no_sum)
    (graph_conv_layer.py:74-103)."""
    ops = ops or _default_ops()
    m, ofn = tfeats.shape[1], tfeats.shape[2]
    if fused:
        tfeats_sum = ops.location_weight_feat_sum(lw, tfeats.contiguous(), nidxs_lens, nidxs_bgs)  # [pn,m,ofn]
    else:
        wtfeats = (lw.unsqueeze(2) * tfeats).reshape(tfeats.shape[0], -1)
        tfeats_sum = ops.neighbor_sum_feat_gather(wtfeats, cidxs, nidxs_lens, nidxs_bgs).reshape(-1, m, ofn)
    pfeats = tfeats_sum * (1.0 / (lw_sum + 1e-6)).unsqueeze(2)
    return pfeats.reshape(pfeats.shape[0], -1) if no_sum else pfeats.sum(1)


def _location_weights(sxyz, pmiu, nidxs_lens, nidxs_bgs, ops):
    lw = torch.exp(sxyz @ pmiu)                                        # [csum,m]
    return lw, ops.location_weight_sum(lw, nidxs_lens, nidxs_bgs)      # [pn,m]


def graph_conv_xyz_feats_impl(xyz, feats, cidxs, nidxs, nidxs_lens, nidxs_bgs, pw, compute_lw=False, lw=None,
                              lw_sum=None, pmiu=None, no_sum=False, fused=True, ops=None):
    """pw [3+ifn,m,ofn] (graph_conv_layer.py:105-124) -> (pfeats, lw, lw_sum)."""
    ops = ops or _default_ops()
    sxyz = ops.neighbor_scatter(xyz, nidxs, nidxs_lens, nidxs_bgs, use_diff=True)
    sfeats = ops.neighbor_scatter(feats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False)
    tfeats = (torch.cat([sxyz, sfeats], 1) @ pw.reshape(pw.shape[0], -1)).reshape(-1, pw.shape[1], pw.shape[2])
    if compute_lw:
        assert lw is None and lw_sum is None and pmiu is not None
        lw, lw_sum = _location_weights(sxyz, pmiu, nidxs_lens, nidxs_bgs, ops)
    else:
        assert lw is not None and lw_sum is not None
    return compute_pfeats(lw, lw_sum, tfeats, nidxs_lens, nidxs_bgs, cidxs, no_sum, fused, ops), lw, lw_sum


def graph_conv_xyz_impl(xyz, cidxs, nidxs, nidxs_lens, nidxs_bgs, pw, compute_lw=False, lw=None, lw_sum=None,
                        pmiu=None, no_sum=False, fused=True, ops=None):
    """pw [3,m,ofn] (graph_conv_layer.py:127-144) -> (pfeats, lw, lw_sum)."""
    ops = ops or _default_ops()
    sxyz = ops.neighbor_scatter(xyz, nidxs, nidxs_lens, nidxs_bgs, use_diff=True)
    tfeats = (sxyz @ pw.reshape(pw.shape[0], -1)).reshape(-1, pw.shape[1], pw.shape[2])
    if compute_lw:
        assert lw is None and lw_sum is None and pmiu is not None
        lw, lw_sum = _location_weights(sxyz, pmiu, nidxs_lens, nidxs_bgs, ops)
    else:
        assert lw is not None and lw_sum is not None
    return compute_pfeats(lw, lw_sum, tfeats, nidxs_lens, nidxs_bgs, cidxs, no_sum, fused, ops), lw, lw_sum


def graph_conv_feats_impl(feats, cidxs, nidxs, nidxs_lens, nidxs_bgs, pw, lw, lw_sum, no_sum=False, fused=True,
                          ops=None):
    """pw [ifn,m,ofn]: project at the points, then scatter (graph_conv_layer.py:147-156)."""
    ops = ops or _default_ops()
    tfeats = ops.neighbor_scatter(feats @ pw.reshape(pw.shape[0], -1), nidxs, nidxs_lens, nidxs_bgs, use_diff=False)
    tfeats = tfeats.reshape(-1, pw.shape[1], pw.shape[2])
    return compute_pfeats(lw, lw_sum, tfeats, nidxs_lens, nidxs_bgs, cidxs, no_sum, fused, ops)


def graph_diff_conv_feats_impl(feats, cidxs, nidxs, nidxs_lens, nidxs_bgs, pw, lw, lw_sum, no_sum=False, fused=True,
                               ops=None):
    """difference features scattered first, then projected per edge (graph_conv_layer.py:159-165)."""
    ops = ops or _default_ops()
    tfeats = ops.neighbor_scatter(feats, nidxs, nidxs_lens, nidxs_bgs, use_diff=True) @ pw.reshape(pw.shape[0], -1)
    tfeats = tfeats.reshape(-1, pw.shape[1], pw.shape[2])
    return compute_pfeats(lw, lw_sum, tfeats, nidxs_lens, nidxs_bgs, cidxs, no_sum, fused, ops)


class _DenseFC(nn.Module):
    """`for fd in fc_dims: x = concat([relu(fc(x)), x])` — the loop every edge layer of the reference opens with."""

    def __init__(self, in_dim, fc_dims):
        super().__init__()
        self.fcs = nn.ModuleList()
        d = in_dim
        for fd in fc_dims:
            self.fcs.append(nn.Linear(d, fd))
            d += fd
        self.out_dim = d

    def forward(self, x):
        for fc in self.fcs:
            x = torch.cat([F.relu(fc(x)), x], 1)
        return x


def _mean_gather(edge_feats, cidxs, nidxs_lens, nidxs_bgs, ops, eps=0.0):
    inv = ((1.0 + eps) / (nidxs_lens.to(torch.float32) + eps)).unsqueeze(1)
    return inv * ops.neighbor_sum_feat_gather(edge_feats, cidxs, nidxs_lens, nidxs_bgs)


class GraphConvEdge(nn.Module):
    """graph_conv_edge (graph_conv_layer.py:573-601): an MLP on [feats_j - feats_i | sxyz] emits a per-edge
    [ifn, ofn] weight matrix, applied to feats_j; mean over the neighbourhood; bias; activation."""

    def __init__(self, ifn, fc_dims, ofn, activation=F.relu, ops=None):
        super().__init__()
        self.ifn, self.ofn, self.activation, self.ops = ifn, ofn, activation, ops
        self.mlp = _DenseFC(ifn + 3, fc_dims)
        self.fc_ew = nn.Linear(self.mlp.out_dim, ifn * ofn)
        self.bias = nn.Parameter(torch.zeros(ofn))

    def forward(self, sxyzs, feats, nidxs, nidxs_lens, nidxs_bgs, cidxs):
        ops = self.ops or _default_ops()
        sfeats = torch.cat([ops.neighbor_scatter(feats, nidxs, nidxs_lens, nidxs_bgs, use_diff=True), sxyzs], 1)
        ew = self.fc_ew(self.mlp(sfeats)).reshape(-1, self.ifn, self.ofn)
        nf = ops.neighbor_scatter(feats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False)
        ef = torch.bmm(nf.unsqueeze(1), ew).squeeze(1)                                 # [en,ofn]
        out = _mean_gather(ef, cidxs, nidxs_lens, nidxs_bgs, ops) + self.bias
        return self.activation(out) if self.activation is not None else out


class GraphConvEdgeXyz(nn.Module):
    """graph_conv_edge_xyz (graph_conv_layer.py:603-627): same with the coordinates as the only input (ifn = 3)."""

    def __init__(self, fc_dims, ofn, activation=F.relu, ops=None):
        super().__init__()
        self.ofn, self.activation, self.ops = ofn, activation, ops
        self.mlp = _DenseFC(3, fc_dims)
        self.fc_ew = nn.Linear(self.mlp.out_dim, 3 * ofn)
        self.bias = nn.Parameter(torch.zeros(ofn))

    def forward(self, sxyzs, nidxs, nidxs_lens, nidxs_bgs, cidxs):
        ops = self.ops or _default_ops()
        ew = self.fc_ew(self.mlp(sxyzs)).reshape(-1, 3, self.ofn)
        ef = torch.bmm(sxyzs.unsqueeze(1), ew).squeeze(1)
        out = _mean_gather(ef, cidxs, nidxs_lens, nidxs_bgs, ops) + self.bias
        return self.activation(out) if self.activation is not None else out


class EcdFeats(nn.Module):
    """ecd_feats (model_pgnet.py:4-32): phi-MLP on [feats_j - feats_i | edge_coord] -> tanh gate [en, ifn];
    gate * feats_j -> g-MLP (dense concat) -> (1+eps)/(len+eps) * sum over the neighbourhood -> fc + relu."""

    def __init__(self, input_dim, phi_dims, g_dims, output_dim, ops=None):
        super().__init__()
        self.ops = ops
        self.phi = _DenseFC(input_dim + 3, phi_dims)
        self.fc_ew = nn.Linear(self.phi.out_dim, input_dim)
        self.g = _DenseFC(input_dim, g_dims)
        self.fc_out = nn.Linear(self.g.out_dim, output_dim)

    def forward(self, edge_coord, point_feats, nidxs, nlens, nbegs, ncens):
        ops = self.ops or _default_ops()
        d = ops.neighbor_scatter(point_feats, nidxs, nlens, nbegs, use_diff=True)
        gate = torch.tanh(self.fc_ew(self.phi(torch.cat([d, edge_coord], 1))))
        ef = self.g(gate * ops.neighbor_scatter(point_feats, nidxs, nlens, nbegs, use_diff=False))
        return F.relu(self.fc_out(_mean_gather(ef, ncens, nlens, nbegs, ops, eps=1e-3)))

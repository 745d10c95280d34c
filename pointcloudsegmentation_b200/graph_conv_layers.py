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

    def __init__(self, input_dim, phi_dims, g_dims, output_dim, ops=None, split=True):
        super().__init__()
        self.ops = ops
        self.split = split
        self.input_dim = input_dim
        self.phi = _DenseFC(input_dim + 3, phi_dims)
        self.fc_ew = nn.Linear(self.phi.out_dim, input_dim)
        self.g = _DenseFC(input_dim, g_dims)
        self.fc_out = nn.Linear(self.g.out_dim, output_dim)

    def _gate_split(self, edge_coord, point_feats, nidxs, nlens, nbegs, ncens):
        """The phi-MLP + gate with the linear layers distributed over the dense concat (same function, same
        parameters; only the fp32 summation order differs).  Every layer's weight is [W_h | W_d | W_x] over
        [h_{k-1} .. h_1 | f_j - f_i | xyz], so the d- and xyz-parts of ALL layers are one point-level GEMM
        P = f [W_d]^T followed by one pass over the edges, u[e] = b - P[i] + P[j] + W_x xyz_e (fused_ops.edge_combine):
        neither the difference tensor [en, C] nor the ever wider concats [en, C+3+..] are materialised."""
        from . import fused_ops
        from .network import edge_addmm, linear_tall

        C = self.input_dim
        layers = list(self.phi.fcs) + [self.fc_ew]
        outs = [l.out_features for l in layers]
        nh = [sum(outs[:k][::-1]) for k in range(len(layers))]  # width of the h-part of layer k's input
        Wd = torch.cat([l.weight[:, n:n + C] for l, n in zip(layers, nh)], 0)
        Wx = torch.cat([l.weight[:, n + C:] for l, n in zip(layers, nh)], 0)
        b = torch.cat([l.bias for l in layers], 0)
        P = linear_tall(point_feats, Wd)
        u = fused_ops.edge_combine(torch.cat([b - P, P], 1), edge_coord, Wx, nidxs, nlens, nbegs, ncens)
        parts = u.split(outs, dim=1)
        hs = []  # most recent first, the order the reference's concat puts them in
        for l, n, pre in zip(layers[:-1], nh[:-1], parts[:-1]):
            if hs:
                pre = edge_addmm(pre, hs[0] if len(hs) == 1 else torch.cat(hs, 1), l.weight[:, :n])
            hs.insert(0, F.relu(pre))
        return torch.tanh(edge_addmm(parts[-1], hs[0] if len(hs) == 1 else torch.cat(hs, 1), self.fc_ew.weight[:, :nh[-1]]))

    def forward(self, edge_coord, point_feats, nidxs, nlens, nbegs, ncens):
        ops = self.ops or _default_ops()
        H = self.phi.out_dim - 3  # = sum(phi_dims) + input_dim
        if (self.split and self.ops is None and point_feats.is_cuda and len(self.phi.fcs) >= 1
                and (H % 4 == 0) and H <= 256 and not edge_coord.requires_grad):  # the fused pass has no d/d(xyz)
            gate = self._gate_split(edge_coord, point_feats, nidxs, nlens, nbegs, ncens)
        else:
            d = ops.neighbor_scatter(point_feats, nidxs, nlens, nbegs, use_diff=True)
            gate = torch.tanh(self.fc_ew(self.phi(torch.cat([d, edge_coord], 1))))
        ef0 = gate * ops.neighbor_scatter(point_feats, nidxs, nlens, nbegs, use_diff=False)
        if self.split and self.ops is None and point_feats.is_cuda:
            # g-MLP without its concats: the x-parts of all layers are one edge GEMM, and the neighbourhood sum is
            # linear, so [g_k .. g_1 | ef0] is summed piece by piece and concatenated at the POINT level
            from .network import edge_addmm

            layers = list(self.g.fcs)
            outs = [l.out_features for l in layers]
            nh = [sum(outs[:k]) for k in range(len(layers))]
            pre = F.linear(ef0, torch.cat([l.weight[:, n:] for l, n in zip(layers, nh)], 0), torch.cat([l.bias for l in layers], 0))
            hs = []
            for l, n, q in zip(layers, nh, pre.split(outs, dim=1)):
                if hs:
                    q = edge_addmm(q, hs[0] if len(hs) == 1 else torch.cat(hs, 1), l.weight[:, :n])
                hs.insert(0, F.relu(q))
            pooled = torch.cat([_mean_gather(x.contiguous(), ncens, nlens, nbegs, ops, eps=1e-3) for x in hs + [ef0]], 1)
            return F.relu(self.fc_out(pooled))
        ef = self.g(ef0)
        return F.relu(self.fc_out(_mean_gather(ef, ncens, nlens, nbegs, ops, eps=1e-3)))

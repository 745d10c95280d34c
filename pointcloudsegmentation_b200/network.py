"""End-to-end harness for the op layer: ONE reference network restated in plain PyTorch — the S3DIS
graph-pool segmentation tower of train_graph_pool_new.py:50-71:

    points_pooling_two_layers(voxel 0.15 / 0.45, block 3.0)        tf_ops/graph_pooling_layer.py:110-146
    pointnet_13_dilated_embed                                       model_pointnet.py:930-1037
    classifier_v3 (1x1 conv 512 -> 256 -> classes, dropout 0.7)     model.py:762-799
    class-weighted softmax cross-entropy                             train_graph_pool_new.py:61-66

Everything that is a custom op in the reference goes through this package's CUDA ops (via the layer
namespace `L`, default = graph_layers); the `fully_connected` / concat glue is stock PyTorch, as it is
stock TensorFlow upstream.  Several independent blocks ride in one block-diagonal batch
(`block_ids`): every custom op is local to a block, and this network has no cross-point reduction
outside them, so the batched result equals the per-block results concatenated.
"""
from __future__ import annotations

import types

import torch
import torch.nn as nn
import torch.nn.functional as F


def tn_tall(a, b, chunk=8192):
    from .fused_ops import tn_tall as f  # deferred: fused_ops loads the CUDA library

    return f(a, b, chunk)


def _tc_ok(x, n_out):
    """Tall CUDA activations go to the tcgen05 linear kernel (fused_ops.linear_tc); short or narrow
    products stay on the stock GEMM (a persistent 148-CTA launch does not pay below ~8k rows / 32 columns)."""
    return x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.shape[0] >= 8192 and n_out >= 32 and x.shape[1] >= 8


def linear_fwd(x, w, b=None, addend=None):
    """x @ w^T (+ b) (+ addend) without autograd: tensor cores for tall x, stock otherwise."""
    if _tc_ok(x, w.shape[0]):
        from .fused_ops import linear_tc

        n = w.shape[0]
        out = None
        if n % 4:  # rows on a 16-byte pitch (a view of a padded buffer): the kernel then stores 16 bytes at a time
            out = torch.empty((x.shape[0], (n + 3) // 4 * 4), dtype=x.dtype, device=x.device)[:, :n]
        return linear_tc(x, w, b, addend, out=out)
    if addend is not None:
        y = torch.addmm(addend, x, w.t())
        return y + b if b is not None else y
    return F.linear(x, w, b)


def linear_wgrad(g, x, with_bias, chunk=2048):
    """(g^T @ x, column sums of g or None) without autograd: tcgen05 kernel for tall operands (the bias gradient
    rides along as an extra column), slab GEMM + column-sum kernel otherwise."""
    if (g.is_cuda and g.dtype == torch.float32 and g.shape[0] >= 8192 and g.shape[1] >= 16 and x.shape[1] >= 8
            and USE_TC_WGRAD):
        from .fused_ops import linear_tc_wgrad

        return linear_tc_wgrad(g, x, with_bias)
    return tn_tall(g, x, chunk), (column_sum(g) if with_bias else None)


USE_TC_WGRAD = True


def column_sum(x):
    if not x.is_cuda:  # the CPU reference network of the tests runs this module's layers on CPU tensors
        return x.sum(0)
    from .fused_ops import column_sum as f

    return f(x)

# class weights of train_graph_pool_new.py:44-47
S3DIS_TRAIN_WEIGHTS = [2.80089331, 2.92693353, 2.94871211, 5.12748384, 5.07317114, 5.18505001, 4.612535,
                       4.83436918, 4.1070838, 5.36530066, 4.64813137, 5.26789713, 3.67803526]


def _ops_csr_transpose(nidxs, pn):
    from . import ops

    return ops.csr_transpose(nidxs, pn)


def _ops_csr_transpose_symmetric(nidxs, nlens, nbegs, ncens, flag):
    from . import ops

    return ops.csr_transpose_symmetric(nidxs, nlens, nbegs, ncens, flag)


def _ops_transpose_verified(nidxs):
    from . import ops

    ops.transpose_verified(nidxs)


def _ops_relu_concat_dropout(a, b, keep_prob, training):
    from . import fused_ops

    return fused_ops.relu_concat_dropout(a, b, keep_prob, training)


def default_layers():
    """The layer namespace the network calls: this package's CUDA ops."""
    from . import graph_layers as G
    from . import neighbor_ops

    return types.SimpleNamespace(
        points_pooling_two_layers=G.points_pooling_two_layers, search_neighborhood=G.search_neighborhood,
        search_neighborhood_range=G.search_neighborhood_range, search_neighborhood_multi=G.search_neighborhood_multi,
        graph_avg_pool=G.graph_avg_pool,
        graph_concat_scatter=G.graph_concat_scatter, graph_pool=G.graph_pool, graph_unpool=G.graph_unpool,
        neighbor_scatter=neighbor_ops.neighbor_scatter, block_begs_from_ids=G.block_begs_from_ids,
        prepare_graph=_ops_csr_transpose, prepare_graph_symmetric=_ops_csr_transpose_symmetric,
        transpose_verified=_ops_transpose_verified,
        relu_concat_dropout=_ops_relu_concat_dropout)


class _EdgeAddmm(torch.autograd.Function):
    """pre + x @ w^T on tall (edge or point) tensors, with the weight gradient computed by `fused_ops.tn_tall`."""

    @staticmethod
    def forward(ctx, pre, x, w, chunk):
        ctx.save_for_backward(x, w)
        ctx.chunk = chunk
        return linear_fwd(x, w, None, pre)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        if g.stride(1) != 1:
            g = g.contiguous()
        gx = linear_fwd(g, w.t()) if ctx.needs_input_grad[1] else None  # edge offsets (sxyz) carry no gradient
        return g, gx, linear_wgrad(g, x, False, ctx.chunk)[0], None


class _SplitConvWeights(torch.autograd.Function):
    """Every layer weight of a pointnet_conv MLP is [W_h | W_centre | W_neighbour | W_xyz] column-wise
    (h = outputs of the earlier layers, newest first).  -> (Wcn = [Wc; Wn] [2H,C], Wx [H,3], [bias; 0] [2H],
    W_h of layers 2.. : W2 [B,A], W3 [C,B+A], Wo [O,C+B+A]), rows stacked over the layers.  One autograd
    node instead of ~20 slice / cat nodes per convolution."""

    @staticmethod
    def forward(ctx, C, *params):
        ws, bs = params[0::2], params[1::2]
        dims = [w.shape[0] for w in ws]
        nh = [sum(dims[:k]) for k in range(len(ws))]
        ctx.meta = (C, dims, nh)
        Wcn = torch.cat([w[:, n:n + C] for w, n in zip(ws, nh)] + [w[:, n + C:n + 2 * C] for w, n in zip(ws, nh)], 0)
        Wx = torch.cat([w[:, n + 2 * C:n + 2 * C + 3] for w, n in zip(ws, nh)], 0)
        bias2 = torch.cat(list(bs) + [bs[0].new_zeros(sum(dims))], 0)
        return (Wcn, Wx, bias2, *[w[:, :n].contiguous() for w, n in zip(ws[1:], nh[1:])])

    @staticmethod
    def backward(ctx, dWcn, dWx, dbias2, *dwh):
        C, dims, nh = ctx.meta
        ref = next(g for g in (dWcn, dWx, dbias2, *dwh) if g is not None)
        z = lambda g, shape: g if g is not None else ref.new_zeros(shape)  # noqa: E731
        H = sum(dims)
        dWcn, dWx, dbias2 = z(dWcn, (2 * H, C)), z(dWx, (H, 3)), z(dbias2, (2 * H,))
        out = [None]
        off = 0
        for k, d in enumerate(dims):
            parts = [dWcn[off:off + d], dWcn[H + off:H + off + d], dWx[off:off + d]]
            if k > 0:
                parts.insert(0, z(dwh[k - 1], (d, nh[k])))
            out += [torch.cat(parts, 1), dbias2[off:off + d]]
            off += d
        return tuple(out)


class _LinearTall(torch.autograd.Function):
    """F.linear for tall activations ([pn ~ 10^5 rows]): same forward; the bias gradient is a column sum of
    a tall matrix (fused_ops.column_sum: an HBM stream instead of the stock reduction), and with `slab` the
    weight gradient g^T x (an [out, in] matrix reduced over pn rows) goes through fused_ops.tn_tall, which
    splits the long reduction into slabs — for <= 256 output columns the stock GEMM puts a handful of
    64x64 tiles on 148 SMs."""

    @staticmethod
    def forward(ctx, x, w, b, slab):
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        ctx.slab = slab
        return linear_fwd(x, w, b)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        if g.stride(1) != 1:  # (a padded row pitch is fine: every consumer below takes a row stride)
            g = g.contiguous()
        dx = linear_fwd(g, w.t()) if ctx.needs_input_grad[0] else None
        dw = db = None
        want_b = ctx.has_bias and ctx.needs_input_grad[2]
        if ctx.needs_input_grad[1] and ctx.slab:
            dw, db = linear_wgrad(g, x, want_b)
        else:
            if ctx.needs_input_grad[1]:
                dw = g.t() @ x
            db = column_sum(g) if want_b else None
        return dx, dw, db, None


def linear_tall(x, w, b=None, slab=True):
    return _LinearTall.apply(x, w, b, slab)


def edge_addmm(pre, x, w, chunk=8192):
    return _EdgeAddmm.apply(pre, x, w, chunk)


class _DenseMLP(nn.Module):
    """fc -> concat([fc_out, in]) repeated, then a linear output layer
    (the loop shared by pointnet_conv / pointnet_pool / mlp, model_pointnet.py:10-80)."""

    def __init__(self, fc_dims, ofn):
        super().__init__()
        self.fcs = nn.ModuleList([nn.LazyLinear(d) for d in fc_dims])
        self.out = nn.LazyLinear(ofn)

    def forward(self, x):
        for fc in self.fcs:
            x = torch.cat([F.relu(fc(x)), x], dim=1)
        return self.out(x)


class PointnetConv(nn.Module):
    """pointnet_conv, model_pointnet.py:10-24: concat-scatter -> edge MLP -> max-gather.

    `split=False` evaluates it literally as the reference does: build x0 = [feats[i] | feats[j] | sxyz]
    per edge ([E, 2C+3]), then the dense-concat MLP on ever wider edge tensors.
    `split=True` (default on the GPU) evaluates the SAME function with the linear layers distributed
    over the concat: every layer's weight is [W_h | W_centre | W_neighbour | W_xyz], so
        x_l @ W^T = h-part @ W_h^T + (feats @ W_centre^T)[i] + (feats @ W_neighbour^T)[j] + sxyz @ W_xyz^T
    and the two feats products are POINT-level GEMMs (k-bar times fewer rows), gathered to the edges
    once for all layers.  Edge tensors shrink from 2C+3+sum(fc) to sum(fc)+ofn columns and the edge
    GEMMs from ~10 kMAC to ~1.5 kMAC per edge.  Same parameters, same state_dict, fp32 throughout;
    only the summation order differs (tests/test_gpu_network.py compares both forms)."""

    def __init__(self, fc_dims, ofn, L, split=True, fused=True):
        super().__init__()
        self.mlp = _DenseMLP(fc_dims, ofn)
        self.L = L
        self.split = split
        self.fused = fused  # edge MLP in one CUDA kernel (fused_ops.dense_mlp3) when the widths allow

    def forward(self, sxyzs, feats, nidxs, nlens, nbegs, ncens):
        layers = list(self.mlp.fcs) + [self.mlp.out]
        if self.split and not any(nn.parameter.is_lazy(l.weight) for l in layers):
            return self._forward_split(layers, sxyzs, feats, nidxs, nlens, nbegs, ncens)
        sfeats = self.L.graph_concat_scatter(feats, nidxs, nlens, nbegs, ncens)
        sfeats = torch.cat([sfeats, sxyzs], dim=1)
        return self.L.graph_pool(self.mlp(sfeats), nlens, nbegs)

    def _forward_split(self, layers, sxyzs, feats, nidxs, nlens, nbegs, ncens):
        L = self.L
        C = feats.shape[1]
        dims = [l.out_features for l in layers]
        nh = [sum(dims[:k]) for k in range(len(layers))]  # width of the h-part of layer k's input
        Wcn, Wx, bias2, *w_h = _SplitConvWeights.apply(C, *[p for l in layers for p in (l.weight, l.bias)])
        pab = linear_tall(feats, Wcn, bias2)  # point level: [pa | pb] = feats [Wc; Wn]^T + [bias; 0]
        H = Wx.shape[0]
        if self.fused and pab.is_cuda and H % 4 == 0 and H <= 256:
            from . import fused_ops

            u = fused_ops.edge_combine(pab, sxyzs, Wx, nidxs, nlens, nbegs, ncens)  # [E, H], one pass
        else:
            u = L.graph_unpool(pab[:, :H], nlens, nbegs, ncens) + L.neighbor_scatter(pab[:, H:], nidxs, nlens, nbegs,
                                                                                    use_diff=False)
            u = edge_addmm(u, sxyzs, Wx)
        if self.fused and u.is_cuda and len(layers) == 4:
            from . import fused_ops

            if fused_ops.dense_mlp3_supported(*dims):
                return fused_ops.dense_mlp3_pool(u, *w_h, nlens, nbegs, ncens)[0]
        parts = u.split(dims, dim=1)  # one cat in backward instead of a zero-filled [E, H] per slice
        hs = []  # most recent first, the order the reference's concat puts them in
        for w, pre in zip([None] + w_h[:-1], parts[:-1]):
            if hs:
                pre = edge_addmm(pre, hs[0] if len(hs) == 1 else torch.cat(hs, 1), w)
            hs.insert(0, F.relu(pre))
        y = edge_addmm(parts[-1], hs[0] if len(hs) == 1 else torch.cat(hs, 1), w_h[-1])
        return L.graph_pool(y, nlens, nbegs)


class PointnetPool(nn.Module):
    """pointnet_pool, model_pointnet.py:56-69 -> (pooled, per-point).

    With `ncens` (the voxel id of every point) and `split=True` the dense-concat MLP is evaluated like
    PointnetConv's: every layer's weight is [W_h | W_x] over the concat [h_{k-1} .. h_1 | x], so all the
    x-products are ONE point-level GEMM  u = x [W1x; W2x; W3x; Wox]^T + [b1; b2; b3; bo]  and the rest is
    the fused MLP + segmented max (fused_ops.dense_mlp3_pool) over the voxel segments.  The per-point
    output is not materialised in that form (returned as None): the tower only uses the pooled one."""

    def __init__(self, fc_dims, ofn, L, split=True):
        super().__init__()
        self.mlp = _DenseMLP(fc_dims, ofn)
        self.L = L
        self.split = split

    def forward(self, xyzs, feats, nlens, nbegs, ncens=None):
        x = torch.cat([xyzs, feats], dim=1)
        layers = list(self.mlp.fcs) + [self.mlp.out]
        if (self.split and ncens is not None and x.is_cuda and len(layers) == 4
                and not any(nn.parameter.is_lazy(l.weight) for l in layers)):
            from . import fused_ops

            dims = [l.out_features for l in layers]
            if fused_ops.dense_mlp3_supported(*dims):
                nh = [sum(dims[:k]) for k in range(4)]
                wx = torch.cat([l.weight[:, n:] for l, n in zip(layers, nh)], 0)
                bias = torch.cat([l.bias for l in layers], 0)
                w_h = [l.weight[:, :n].contiguous() for l, n in zip(layers[1:], nh[1:])]
                u = linear_tall(x, wx, bias)
                return fused_ops.dense_mlp3_pool(u, *w_h, nlens, nbegs, ncens)[0], None
        s = self.mlp(x)
        return self.L.graph_pool(s, nlens, nbegs), s


class _Embed(nn.Module):
    """fc_embed, model_pointnet.py:101-104 (leaky_relu, TF default alpha 0.2)."""

    def __init__(self, dim):
        super().__init__()
        self.fc = nn.LazyLinear(dim)

    def forward(self, x):
        if nn.parameter.is_lazy(self.fc.weight):
            return F.leaky_relu(self.fc(x), 0.2)
        return F.leaky_relu(linear_tall(x, self.fc.weight, self.fc.bias), 0.2)


# (level, r_min or None, r_max, rescale, [(embed_dim or None, fc_dims, ofn), ...]) — model_pointnet.py:935-1021
_STAGES = [
    (0, None, 0.15, 0.15, [(None, [8, 8, 16], 32)]),
    (0, 0.15, 0.2, 0.15, [(None, [8, 8, 16], 32)]),
    (0, 0.1, 0.15, 0.15, [(None, [8, 8, 16], 32)]),
    (0, None, 0.1, 0.15, [(32, [8, 8, 16], 32)]),
    (1, None, 0.45, 0.45, [(64, [16, 16, 32], 64)]),
    (1, 0.45, 0.6, 0.45, [(48, [16, 16, 16], 48), (48, [16, 16, 16], 48)]),
    (1, 0.3, 0.45, 0.45, [(64, [16, 16, 16], 48), (64, [16, 16, 16], 48)]),
    (1, None, 0.3, 0.45, [(96, [16, 16, 16], 48), (96, [16, 16, 16], 48)]),
    (2, None, 0.9, 0.9, [(128, [16, 16, 32], 64), (128, [16, 16, 32], 64)]),
]


class Pointnet13DilatedEmbed(nn.Module):
    """pointnet_13_dilated_embed (model_pointnet.py:930-1037) + classifier_v3 + weighted CE."""

    # the tables a subclass replaces (Semantic3DDense below): graphs and convolutions, the two pooling MLPs, the
    # global MLP, the two voxel sizes and the block size of points_pooling_two_layers
    STAGES = _STAGES
    POOLS = (([8, 8, 16], 32), ([16, 16, 16], 48))
    GLOBAL = ([32, 32, 48], 128)
    VOXELS = (0.15, 0.45)
    BLOCK_SIZE = 3.0

    def __init__(self, num_classes=13, L=None, class_weights=None, dropout_keep=0.7, split=True):
        super().__init__()
        self.L = L or default_layers()
        self.groups = nn.ModuleList()
        for _, _, _, _, convs in self.STAGES:
            g = nn.ModuleList()
            for embed, fc_dims, ofn in convs:
                g.append(nn.ModuleList([_Embed(embed) if embed else nn.Identity(), PointnetConv(fc_dims, ofn, self.L, split)]))
            self.groups.append(g)
        self.pool0 = PointnetPool(*self.POOLS[0], self.L, split)
        self.pool1 = PointnetPool(*self.POOLS[1], self.L, split)
        self.global_mlp = _DenseMLP(*self.GLOBAL)
        self.class_mlp1 = nn.LazyLinear(512)
        self.class_mlp2 = nn.LazyLinear(256)
        self.class_mlp3 = nn.LazyLinear(num_classes)
        self.dropout_keep = dropout_keep
        self.split = split
        w = torch.tensor(class_weights if class_weights is not None else S3DIS_TRAIN_WEIGHTS[:num_classes],
                         dtype=torch.float32)
        self.register_buffer("class_weights", w)

    def build_graphs(self, xyzs, block_begs):
        """The nine neighbourhood graphs of the tower (model_pointnet.py:935-1021): they depend on the
        coordinates only.  Built together, before any feature work, so that the host syncs of the searches
        (csum comes back to size the index tensors) do not interrupt the feature kernels."""
        L = self.L
        graphs = []
        want_t = self.training and hasattr(L, "prepare_graph")
        # radius / annulus graphs are symmetric with ascending rows: their transpose is one binary search per
        # edge (ops.csr_transpose_symmetric); a miss raises below, after the last graph (one host sync)
        flags = (torch.zeros(len(self.STAGES), dtype=torch.int32, device=xyzs[0].device)
                 if (want_t and hasattr(L, "prepare_graph_symmetric") and xyzs[0].is_cuda) else None)
        found = {}
        if hasattr(L, "search_neighborhood_multi") and xyzs[0].is_cuda:
            # all graphs of a level from one traversal of its points, one sizing read per level
            for level in sorted({s[0] for s in self.STAGES}):
                gis = [gi for gi, s in enumerate(self.STAGES) if s[0] == level]
                bb, mb = block_begs[level]
                res = L.search_neighborhood_multi(xyzs[level], [(self.STAGES[gi][1], self.STAGES[gi][2]) for gi in gis],
                                                  block_begs=bb, max_block_pn=mb)
                found.update(zip(gis, res))
        for gi, (level, rmin, rmax, rescale, _) in enumerate(self.STAGES):
            bb, mb = block_begs[level]
            if gi in found:
                g = found[gi]
            elif rmin is None:
                g = L.search_neighborhood(xyzs[level], rmax, block_begs=bb, max_block_pn=mb)
            else:
                g = L.search_neighborhood_range(xyzs[level], rmin, rmax, block_begs=bb, max_block_pn=mb)
            nidxs, nlens, nbegs, ncens = g
            sxyzs = L.neighbor_scatter(xyzs[level], nidxs, nlens, nbegs, use_diff=True) / rescale
            # the transpose CSR the backward (NeighborGather) needs: built here, with the graph, so that a
            # look-ahead prepare() takes it off the training stream too (training mode only: inference has no
            # backward); kept in the tuple so its owner can
            # hand it from stream to stream (it is also cached on `nidxs`, where the op looks for it)
            if flags is not None:
                extra = L.prepare_graph_symmetric(nidxs, nlens, nbegs, ncens, flags[gi])
            else:
                extra = L.prepare_graph(nidxs, nlens.shape[0]) if want_t else None
            graphs.append((sxyzs, nidxs, nlens, nbegs, ncens, extra))
        if flags is not None:
            if int(flags.max().item()):
                raise RuntimeError("build_graphs: a search returned a non-symmetric graph (csr_transpose_symmetric)")
            if hasattr(L, "transpose_verified"):
                for g in graphs:
                    L.transpose_verified(g[1])  # the cached transposes are now trusted without a per-graph read
        return graphs

    def prepare(self, xyzs, feats, labels, block_ids=None, num_blocks=1):
        """Everything that depends on the input points only (no parameter is read): the two-level voxel
        pooling and the neighbourhood graphs.  A training loop may run it one batch ahead, on a second
        stream, like a data loader (bench.py does): forward(prepared=...) then contains no host sync."""
        L = self.L
        with torch.no_grad():
            if block_ids is None:
                xs, dxs, f, lab, vl, vb, vc = L.points_pooling_two_layers(xyzs, feats, labels, *self.VOXELS, self.BLOCK_SIZE)
                bbegs = [(None, None)] * 3
            else:
                xs, dxs, f, lab, vl, vb, vc, bids = L.points_pooling_two_layers(xyzs, feats, labels, *self.VOXELS, self.BLOCK_SIZE,
                                                                               block_ids=block_ids)
                bbegs = []
                for b in bids:
                    bb = L.block_begs_from_ids(b, num_blocks)
                    bbegs.append((bb, int((bb[1:] - bb[:-1]).max())))
            graphs = self.build_graphs(xs, bbegs)
        return xs, dxs, f, lab, vl, vb, vc, graphs

    def features(self, xyzs, dxyzs, feats, vlens, vbegs, vcens, graphs):
        L = self.L
        feats1 = L.graph_avg_pool(feats, vlens[0], vbegs[0], vcens[0])
        feats2 = L.graph_avg_pool(feats1, vlens[1], vbegs[1], vcens[1])
        stage_out = {}
        cur_level = 0
        for (level, _, _, _, _), group, graph in zip(self.STAGES, self.groups, graphs):
            if level != cur_level:  # stage boundary: pool to the next level
                if cur_level == 0:
                    stage_out[0] = feats
                    vpool, _ = self.pool0(dxyzs[0], feats, vlens[0], vbegs[0], vcens[0])
                    feats = torch.cat([feats1, L.graph_pool(feats, vlens[0], vbegs[0]), vpool], dim=1)
                else:
                    stage_out[1] = feats
                    vpool, _ = self.pool1(dxyzs[1], feats, vlens[1], vbegs[1], vcens[1])
                    feats = torch.cat([feats2, L.graph_pool(feats, vlens[1], vbegs[1]), vpool], dim=1)
                cur_level = level
            for embed, conv in group:
                before = feats
                feats = torch.cat([feats, conv(graph[0], embed(feats), *graph[1:5])], dim=1)
        stage_out[2] = feats
        # model_pointnet.py:1022-1028: feats_stage2 takes feats12's output, but the global MLP is fed
        # concat([xyzs[2], feats]) with `feats` as it stood BEFORE feats12 was appended (633 + 3 columns)
        fc2 = self.global_mlp(torch.cat([xyzs[2], before], dim=1))
        lf2 = torch.cat([stage_out[2], fc2], dim=1)
        return lf2, stage_out[1], stage_out[0]

    def upsample(self, lf2, stage1, stage0, vlens, vbegs, vcens):
        """The reference's per-point feature lf0 = [unpool0([unpool1(lf2) | stage1]) | stage0]
        (model_pointnet.py:1029-1035)."""
        L = self.L
        lf1 = torch.cat([L.graph_unpool(lf2, vlens[1], vbegs[1], vcens[1]), stage1], dim=1)
        return torch.cat([L.graph_unpool(lf1, vlens[0], vbegs[0], vcens[0]), stage0], dim=1)

    def class_mlp1_split(self, lf2, stage1, stage0, vlens, vbegs, vcens):
        """class_mlp1(lf0) without ever building lf0: unpool is a row copy, so it commutes with the
        linear layer, and class_mlp1's weight splits over the concat as [W_lf2 | W_stage1 | W_stage0]:
            class_mlp1(lf0) = unpool0(unpool1(lf2 W_lf2^T) + stage1 W_stage1^T) + stage0 W_stage0^T + b.
        The 1474-wide GEMM then runs over level-2 / level-1 rows (40x / 6x fewer than level 0)."""
        L = self.L
        n2, n1 = lf2.shape[1], stage1.shape[1]
        W, b = self.class_mlp1.weight, self.class_mlp1.bias
        z2 = F.linear(lf2, W[:, :n2], b)  # the bias rides up through the two row copies
        z1 = edge_addmm(L.graph_unpool(z2, vlens[1], vbegs[1], vcens[1]), stage1, W[:, n2:n2 + n1], 2048)
        return edge_addmm(L.graph_unpool(z1, vlens[0], vbegs[0], vcens[0]), stage0, W[:, n2 + n1:], 2048)

    def classify(self, z0, pfeats):
        """classifier_v3 (model.py:762-799) from class_mlp1's pre-activation z0 on."""
        p = 1.0 - self.dropout_keep
        if hasattr(self.L, "relu_concat_dropout") and z0.is_cuda and not nn.parameter.is_lazy(self.class_mlp2.weight):
            # relu + concat + dropout in one pass each way, mask regenerated in the backward (csrc/relu_concat_dropout.cu)
            # The concatenated rows (512+131, 256+131 floats) are padded with zero columns to a multiple of 16 bytes —
            # and the weights with matching zero columns — so that the tensor-core layers read and write them with
            # 16-byte accesses (linear_tc: 289 -> 213 us per [133 k,643] -> 256); same products, same gradients.
            w2, w3 = self.class_mlp2.weight, self.class_mlp3.weight
            pad2, pad3 = (-w2.shape[1]) % 4, (-w3.shape[1]) % 4
            pf2 = F.pad(pfeats, (0, pad2)) if pad2 else pfeats
            pf3 = pf2 if pad3 == pad2 else (F.pad(pfeats, (0, pad3)) if pad3 else pfeats)
            x = self.L.relu_concat_dropout(z0, pf2, self.dropout_keep, self.training)
            h = linear_tall(x, F.pad(w2, (0, pad2)) if pad2 else w2, self.class_mlp2.bias, True)
            x = self.L.relu_concat_dropout(h, pf3, self.dropout_keep, self.training)
            return linear_tall(x, F.pad(w3, (0, pad3)) if pad3 else w3, self.class_mlp3.bias, True)
        x = torch.cat([F.relu(z0), pfeats], dim=1)
        x = F.dropout(x, p, self.training)
        lazy = nn.parameter.is_lazy(self.class_mlp2.weight)
        lin = (lambda m, t, slab: m(t)) if lazy else (lambda m, t, slab: linear_tall(t, m.weight, m.bias, slab))
        x = torch.cat([F.relu(lin(self.class_mlp2, x, True)), pfeats], dim=1)
        x = F.dropout(x, p, self.training)
        return lin(self.class_mlp3, x, True)

    def logits_of(self, prepared):
        xs, dxs, f, lab, vl, vb, vc, graphs = prepared
        lf2, stage1, stage0 = self.features(xs, dxs, f, vl, vb, vc, graphs)
        if self.split and not nn.parameter.is_lazy(self.class_mlp1.weight):
            z0 = self.class_mlp1_split(lf2, stage1, stage0, vl, vb, vc)
        else:
            z0 = self.class_mlp1(self.upsample(lf2, stage1, stage0, vl, vb, vc))
        return self.classify(z0, stage0)

    def forward(self, xyzs=None, feats=None, labels=None, block_ids=None, num_blocks=1, prepared=None):
        """xyzs [pn,3] (block-normalised), feats [pn,C], labels [pn] -> (loss, logits, permuted labels);
        or prepared = self.prepare(...) of the same arguments."""
        if prepared is None:
            prepared = self.prepare(xyzs, feats, labels, block_ids, num_blocks)
        lab = prepared[3]
        logits = self.logits_of(prepared)
        w = self.class_weights[lab]
        ce = F.cross_entropy(logits, lab, reduction="none")
        loss = (ce * w).sum() / torch.count_nonzero(w).clamp(min=1)  # tf.losses SUM_BY_NONZERO_WEIGHTS
        return loss, logits, lab


# ---- Semantic3D dense model (BASELINE configs[4]) ---------------------------------------------------------------
# pointnet_13_dilate_embed_semantic3d, model_pointnet_semantic3d.py:325-437
_STAGES_SEMANTIC3D = [
    (0, None, 0.3, 0.3, [(32, [8, 8, 16], 32)]),
    (0, 0.3, 0.4, 0.3, [(32, [8, 8, 16], 32)]),
    (0, 0.2, 0.3, 0.3, [(32, [8, 8, 16], 32)]),
    (0, None, 0.2, 0.3, [(32, [8, 8, 16], 32)]),
    (1, None, 1.25, 1.25, [(64, [16, 16, 32], 64)]),
    (1, 1.25, 1.6, 1.25, [(64, [12, 12, 24], 48), (64, [12, 12, 24], 48)]),
    (1, 0.9, 1.25, 1.25, [(64, [12, 12, 24], 48), (64, [12, 12, 24], 48)]),
    (1, None, 0.9, 1.25, [(64, [12, 12, 24], 48), (64, [12, 12, 24], 48)]),
    (2, None, 4.0, 4.0, [(128, [32, 32, 32], 96), (160, [32, 32, 64], 128)]),
]
# train_gpn_semantic3d_dense.py:47 (label 0 = unlabelled: weight 0)
SEMANTIC3D_TRAIN_WEIGHTS = [0.0, 3.5545287, 3.211912, 3.0997152, 4.7834935, 2.7914784, 4.982284, 5.260409, 5.321266]


class Semantic3DDense(Pointnet13DilatedEmbed):
    """tower_loss of train_gpn_semantic3d_dense.py:49-85: dense_feats (a point-net over the DENSE cloud's radius
    neighbourhood of every sampled centre, model_pointnet_semantic3d.py:307-323) -> points_pooling_two_layers(0.45, 1.5,
    10.0) -> pointnet_13_dilate_embed_semantic3d -> classifier_v3 -> weighted CE with label 0 masked."""

    STAGES = _STAGES_SEMANTIC3D
    POOLS = (([8, 8, 16], 32), ([16, 16, 32], 64))
    GLOBAL = ([32, 32, 64], 128)
    VOXELS = (0.45, 1.5)
    BLOCK_SIZE = 10.0

    def __init__(self, num_classes=8, L=None, dropout_keep=0.7, split=True):
        super().__init__(num_classes, L, class_weights=SEMANTIC3D_TRAIN_WEIGHTS[:num_classes + 1], dropout_keep=dropout_keep,
                         split=split)
        self.dense_mlp = _DenseMLP([16, 16, 16], 48)

    def dense_feats(self, xyzs, feats, lbls, idxs, nidxs, nlens, nbegs, ncens):
        """model_pointnet_semantic3d.py:307-323 -> (cxyzs [M,3], cfeats [M,48+C], clbls [M])."""
        L = self.L
        il = idxs.long()
        cxyzs, cfeats, clbls = xyzs[il], feats[il], lbls[il]
        nl = nidxs.long()
        dxyzs = xyzs[nl] - L.graph_unpool(cxyzs, nlens, nbegs, ncens)
        pfeats = torch.cat([dxyzs, L.graph_unpool(cfeats, nlens, nbegs, ncens), feats[nl]], dim=1)
        pfeats = L.graph_pool(self.dense_mlp(pfeats), nlens, nbegs)
        return cxyzs, torch.cat([pfeats, cfeats], dim=1), clbls

    def prepare_points(self, xyzs, lbls, idxs):
        """The parameter-free part of a step (a loader may run it one tile ahead, like Pointnet13DilatedEmbed.prepare):
        two-level voxel pooling of the centres and the nine graphs.  The pooling's row permutation is recovered by
        sending the centre index through it as a feature column (exact below 2^24 centres), so the learned features
        can be permuted later with one gather.  -> (xs, dxs, perm, lab, vl, vb, vc, graphs)."""
        il = idxs.long()
        cx, cl = xyzs[il].contiguous(), lbls[il]
        m = cx.shape[0]
        assert m < (1 << 24), "centre index does not fit a float32 feature column"
        with torch.no_grad():
            tag = torch.arange(m, dtype=torch.float32, device=cx.device).unsqueeze(1)
            xs, dxs, f, lab, vl, vb, vc = self.L.points_pooling_two_layers(cx, tag, cl, *self.VOXELS, self.BLOCK_SIZE)
            perm = f[:, 0].long()
            graphs = self.build_graphs(xs, [(None, None)] * 3)
        return xs, dxs, perm, lab, vl, vb, vc, graphs

    def prepare_dense(self, xyzs, feats, lbls, idxs, nidxs, nlens, nbegs, ncens, points=None):
        if points is None:
            points = self.prepare_points(xyzs, lbls, idxs)
        xs, dxs, perm, lab, vl, vb, vc, graphs = points
        _, cf, _ = self.dense_feats(xyzs, feats, lbls, idxs, nidxs, nlens, nbegs, ncens)
        return xs, dxs, cf[perm], lab, vl, vb, vc, graphs

    def forward(self, xyzs, feats, lbls, idxs, nidxs, nlens, nbegs, ncens, points=None):
        """Dense block (xyzs [N,3] block-normalised to [0,10) m in xy, feats [N,4], lbls [N] in 0..num_classes with
        0 = unlabelled), sampled centres idxs [M], CSR of every centre's dense neighbourhood
        -> (loss, logits [M, num_classes], labels in 0..num_classes-1 as the loss sees them, mask of unlabelled).
        `points` = self.prepare_points(xyzs, lbls, idxs) computed ahead."""
        prepared = self.prepare_dense(xyzs, feats, lbls, idxs, nidxs, nlens, nbegs, ncens, points)
        lab = prepared[3]
        logits = self.logits_of(prepared)
        w = self.class_weights[lab]
        unknown = lab == 0
        tgt = torch.where(unknown, torch.ones_like(lab), lab) - 1  # train_gpn_semantic3d_dense.py:75-77
        ce = F.cross_entropy(logits, tgt, reduction="none")
        loss = (ce * w).sum() / torch.count_nonzero(w).clamp(min=1)
        return loss, logits, tgt, unknown

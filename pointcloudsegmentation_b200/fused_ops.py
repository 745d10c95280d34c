"""Fused layer pieces beyond the reference's op set: what the reference leaves to a
chain of stock `fully_connected` / `concat` / `relu` calls on edge tensors, done in one kernel.

`dense_mlp3(u, w2, w3, wo)` is the dense-concat edge MLP of pointnet_conv (model_pointnet.py:10-24)
after network.py's split evaluation:  h1=relu(u1); h2=relu(u2+W2 h1); h3=relu(u3+W3[h2,h1]);
y = uo + Wo[h3,h2,h1]  for  u = [u1|u2|u3|uo]  (csrc/dense_mlp.cu).

`edge_combine(pab, sxyzs, wx, nidxs, nlens, nbegs, ncens)` builds that u in one pass from the
point-level product pab = [pa | pb]: u[e] = pa[centre(e)] + pb[neighbour(e)] + wx sxyz[e]  (csrc/edge_conv.cu).
"""
from __future__ import annotations

import torch

from . import ops
from ._lib import check, lib


def tn_tall(a, b, chunk=8192):
    """a^T @ b for tall-skinny a [E, m], b [E, n] (E ~ 10^6, m, n <= 64): one GEMM with K = E and a
    32x16 output runs on a handful of CTAs; a batched GEMM over E/chunk slabs followed by a small sum
    fills the GPU.  Same products, fp32, deterministic (fixed slab order in the final sum)."""
    E = a.shape[0]
    nfull = E // chunk
    if nfull < 8:
        return a.t() @ b
    body = nfull * chunk
    # column slices of a wider row-major tensor go to cuBLAS as they are (leading dimension = row stride)
    if a.stride(1) != 1:
        a = a.contiguous()
    if b.stride(1) != 1:
        b = b.contiguous()
    out = torch.bmm(a[:body].view(nfull, chunk, a.shape[1]).transpose(1, 2), b[:body].view(nfull, chunk, b.shape[1])).sum(0)
    if body < E:
        out = out + a[body:].t() @ b[body:]
    return out


def column_sum(x, small_stock=True):
    """x.sum(0) for a tall [rows, ...] fp32 CUDA tensor (csrc/column_sum.cu): bias gradients and partial folds.
    Fixed summation order; no autograd (it is used inside backward passes)."""
    if not x.is_cuda or x.dtype != torch.float32:
        raise ValueError("column_sum: must be a float32 CUDA tensor (this library has no CPU path)")
    if small_stock and x.numel() < (1 << 20):  # a few hundred partial rows: one stock launch is as fast as two of ours
        return x.sum(0)
    x = x.contiguous()
    rows = x.shape[0]
    C = x.numel() // rows if rows else int(torch.Size(x.shape[1:]).numel())
    out = torch.empty(x.shape[1:], dtype=torch.float32, device=x.device)
    P = int(lib.pcs_column_sum_partials(rows, C))
    part = torch.empty((P, C), dtype=torch.float32, device=x.device) if P else None
    check(lib.pcs_column_sum(x.data_ptr(), rows, C, part.data_ptr() if P else None, out.data_ptr(), _stream()),
          "column_sum")
    return out


def linear_tc(x, w, bias=None, addend=None, relu=False, out=None, _dbg=0):
    """y = x @ w^T (+ bias) (+ addend) on the tcgen05 tensor cores at fp32 accuracy (csrc/linear_tc.cu: three
    TF32 MMAs per product on a hi/lo split, fp32 accumulators in TMEM).  x [M,K] (row stride free, unit
    column stride), w [N,K] with ANY strides (a transposed view or a column slice costs nothing: the weight is
    re-laid into its shared-memory image first), bias [N], addend [M,N].  No autograd (used inside Functions)."""
    if not (x.is_cuda and w.is_cuda) or x.dtype != torch.float32 or w.dtype != torch.float32:
        raise ValueError("linear_tc: x and w must be float32 CUDA tensors (this library has no CPU path)")
    if x.dim() != 2 or w.dim() != 2 or x.shape[1] != w.shape[1]:
        raise ValueError(f"linear_tc: shapes {tuple(x.shape)} x {tuple(w.shape)}^T do not match")
    if x.stride(1) != 1 or (x.shape[0] > 1 and x.stride(0) < x.shape[1]):
        x = x.contiguous()
    M, K = x.shape
    N = w.shape[0]
    if bias is not None:
        bias = bias.contiguous()
        if bias.shape != (N,):
            raise ValueError("linear_tc: bias must be [N]")
    if addend is not None:
        if addend.shape != (M, N):
            raise ValueError("linear_tc: addend must be [M,N]")
        if addend.stride(1) != 1:
            addend = addend.contiguous()
    y = out if out is not None else torch.empty((M, N), dtype=torch.float32, device=x.device)
    if y.shape != (M, N) or y.stride(1) != 1:
        raise ValueError("linear_tc: out must be [M,N] with unit column stride")
    nbytes = int(lib.pcs_linear_tc_workspace_bytes(N, K))
    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=x.device)
    check(lib.pcs_linear_tc(x.data_ptr(), M, K, x.stride(0) if M > 1 else K, w.data_ptr(), N, w.stride(0), w.stride(1),
                            bias.data_ptr() if bias is not None else None,
                            addend.data_ptr() if addend is not None else None,
                            addend.stride(0) if addend is not None else 0, (1 if relu else 0) | (_dbg << 8),
                            y.data_ptr(), y.stride(0) if M > 1 else N, ws.data_ptr(), nbytes, _stream()), "linear_tc")
    return y


def linear_tc_wgrad(g, x, with_bias=False):
    """(g^T @ x [N,K], column sums of g [N] or None) for tall g [M,N], x [M,K] on the tcgen05 tensor cores
    (csrc/linear_tc_wgrad.cu): both operands are read row-major as they are (MN-major UMMA operands), the M rows are
    split over CTAs and the partials folded in fixed order.  The bias gradient rides along as one extra column."""
    if not (g.is_cuda and x.is_cuda) or g.dtype != torch.float32 or x.dtype != torch.float32:
        raise ValueError("linear_tc_wgrad: g and x must be float32 CUDA tensors (this library has no CPU path)")
    if g.dim() != 2 or x.dim() != 2 or g.shape[0] != x.shape[0] or g.shape[0] == 0:
        raise ValueError(f"linear_tc_wgrad: shapes {tuple(g.shape)}, {tuple(x.shape)} do not match")
    if g.stride(1) != 1:
        g = g.contiguous()
    if x.stride(1) != 1:
        x = x.contiguous()
    M, N = g.shape
    K = x.shape[1]
    wb = 1 if with_bias else 0
    S = int(lib.pcs_linear_tc_wgrad_partials(M, N, K, wb))
    part = torch.empty((S, N, K + wb), dtype=torch.float32, device=g.device)
    check(lib.pcs_linear_tc_wgrad(g.data_ptr(), M, N, g.stride(0) if M > 1 else N, x.data_ptr(), K,
                                  x.stride(0) if M > 1 else K, wb, part.data_ptr(), _stream()), "linear_tc_wgrad")
    out = part[0] if S == 1 else column_sum(part)
    if with_bias:
        return out[:, :K], out[:, K]
    return out, None


def _rcd_check(a, b):
    for t, n in ((a, "a"), (b, "b")):
        if not t.is_cuda or t.dtype != torch.float32 or t.dim() != 2:
            raise ValueError(f"relu_concat_dropout: {n} must be a 2-D float32 CUDA tensor (this library has no CPU path)")
    if a.shape[0] != b.shape[0] or a.shape[1] == 0:
        raise ValueError(f"relu_concat_dropout: shapes {tuple(a.shape)}, {tuple(b.shape)} do not match")


def relu_concat_dropout_forward(a, b, keep_prob=1.0, seed=0):
    """dropout(cat([relu(a), b], 1), keep_prob) in one pass (csrc/relu_concat_dropout.cu); keep_prob >= 1: no
    dropout.  The mask is a pure function of (seed, element), nothing is stored for the backward."""
    _rcd_check(a, b)
    if a.stride(1) != 1:
        a = a.contiguous()
    if b.shape[1] and b.stride(1) != 1:
        b = b.contiguous()
    M, Na = a.shape
    Nb = b.shape[1]
    out = torch.empty((M, Na + Nb), dtype=torch.float32, device=a.device)
    check(lib.pcs_relu_concat_dropout_forward(a.data_ptr(), M, Na, a.stride(0) if M > 1 else Na,
                                              b.data_ptr() if Nb else None, Nb, (b.stride(0) if M > 1 else Nb) if Nb else 0,
                                              float(keep_prob), int(seed), out.data_ptr(), _stream()), "relu_concat_dropout_forward")
    return out


def relu_concat_dropout_backward(g, out, Na, keep_prob=1.0, seed=0, need_b=True):
    """(d a, d b) of relu_concat_dropout_forward from the gradient g of its output and the output itself."""
    if g.shape != out.shape or not g.is_cuda or g.dtype != torch.float32:
        raise ValueError("relu_concat_dropout_backward: g must match out (float32 CUDA)")
    g = g.contiguous()
    M, W = out.shape
    Nb = W - Na
    da = torch.empty((M, Na), dtype=torch.float32, device=g.device)
    db = torch.empty((M, Nb), dtype=torch.float32, device=g.device) if (need_b and Nb) else None
    check(lib.pcs_relu_concat_dropout_backward(g.data_ptr(), out.data_ptr(), M, Na, Nb, float(keep_prob), int(seed),
                                               da.data_ptr(), Na, db.data_ptr() if db is not None else None, Nb,
                                               _stream()), "relu_concat_dropout_backward")
    return da, db


class _ReluConcatDropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, keep_prob, seed):
        out = relu_concat_dropout_forward(a, b, keep_prob, seed)
        ctx.save_for_backward(out)
        ctx.cfg = (a.shape[1], keep_prob, seed)
        return out

    @staticmethod
    def backward(ctx, g):
        (out,) = ctx.saved_tensors
        Na, keep_prob, seed = ctx.cfg
        da, db = relu_concat_dropout_backward(g, out, Na, keep_prob, seed, need_b=ctx.needs_input_grad[1])
        return da if ctx.needs_input_grad[0] else None, db, None, None


def relu_concat_dropout(a, b, keep_prob=1.0, training=True, seed=None):
    """classifier_v3's `dropout(concat([relu(fc), pfeats]), keep_prob)` (model.py:783-793) with autograd.  `seed`
    defaults to a draw from torch's CPU generator mixed with the rank (torch.manual_seed makes a run repeatable; no device sync)."""
    if not training or keep_prob >= 1.0:
        keep_prob, seed = 1.0, 0
    elif seed is None:
        seed = int(torch.randint(0, 2 ** 62, (), dtype=torch.int64))
        if torch.distributed.is_available() and torch.distributed.is_initialized():
            # ranks seeded alike (torch.manual_seed(0) everywhere) must still draw independent masks, like the
            # reference's towers do
            seed ^= (torch.distributed.get_rank() + 1) * 0x9E3779B97F4A7C15 & (2 ** 62 - 1)
    return _ReluConcatDropout.apply(a, b, float(keep_prob), int(seed))


def dense_mlp3_supported(A, B, C, O) -> bool:
    return bool(lib.pcs_dense_mlp3_supported(int(A), int(B), int(C), int(O)))


def _stream():
    # raw handle of the current stream: torch.cuda.current_stream() builds a Stream object (~5 us per op call)
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


class _DenseMLP3(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, w2, w3, wo):
        u, w2, w3, wo = u.contiguous(), w2.contiguous(), w3.contiguous(), wo.contiguous()
        B, A = w2.shape
        C = w3.shape[0]
        O = wo.shape[0]
        rows = u.shape[0]
        if u.shape[1] != A + B + C + O or w3.shape[1] != A + B or wo.shape[1] != A + B + C:
            raise ValueError("dense_mlp3: inconsistent layer widths")
        y = torch.empty((rows, O), dtype=torch.float32, device=u.device)
        check(lib.pcs_dense_mlp3_forward(u.data_ptr(), w2.data_ptr(), w3.data_ptr(), wo.data_ptr(), rows, A, B, C, O,
                                         y.data_ptr(), _stream()), "dense_mlp3_forward")
        ctx.save_for_backward(u, w2, w3, wo)
        return y

    @staticmethod
    def backward(ctx, dy):
        u, w2, w3, wo = ctx.saved_tensors
        B, A = w2.shape
        C, O = w3.shape[0], wo.shape[0]
        rows = u.shape[0]
        dy = dy.contiguous()
        du = torch.empty_like(u)
        hcat = torch.empty((rows, A + B + C), dtype=torch.float32, device=u.device)
        check(lib.pcs_dense_mlp3_backward(u.data_ptr(), dy.data_ptr(), w2.data_ptr(), w3.data_ptr(), wo.data_ptr(),
                                          rows, A, B, C, O, du.data_ptr(), hcat.data_ptr(), _stream()),
              "dense_mlp3_backward")
        dwo = tn_tall(dy, hcat)                                   # [O, C+B+A]
        dw3 = tn_tall(du[:, A + B:A + B + C], hcat[:, C:])        # [C, B+A]
        dw2 = tn_tall(du[:, A:A + B], hcat[:, C + B:])            # [B, A]
        return du, dw2, dw3, dwo


def dense_mlp3(u, w2, w3, wo):
    """u [rows, A+B+C+O] fp32 CUDA; w2 [B,A], w3 [C,B+A], wo [O,C+B+A] -> y [rows, O]; differentiable."""
    if not u.is_cuda:
        raise ValueError("dense_mlp3: must be a CUDA tensor (this library has no CPU path)")
    return _DenseMLP3.apply(u, w2, w3, wo)


class _EdgeCombine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pab, sxyzs, wx, nidxs, nlens, nbegs, ncens):
        pab, sxyzs, wx = pab.contiguous(), sxyzs.contiguous(), wx.contiguous()
        csum, H = nidxs.shape[0], pab.shape[1] // 2
        if pab.shape[1] != 2 * H or wx.shape != (H, 3) or sxyzs.shape != (csum, 3) or ncens.shape[0] != csum:
            raise ValueError("edge_combine: inconsistent shapes")
        u = torch.empty((csum, H), dtype=torch.float32, device=pab.device)
        check(lib.pcs_edge_combine_forward(pab.data_ptr(), pab.data_ptr() + 4 * H, sxyzs.data_ptr(), wx.data_ptr(),
                                           ncens.data_ptr(), nidxs.data_ptr(), nlens.data_ptr(), nbegs.data_ptr(),
                                           pab.shape[0], csum, H, 2 * H, u.data_ptr(), _stream()),
              "edge_combine_forward")
        ctx.save_for_backward(sxyzs, nidxs, nlens, nbegs, ncens)
        return u

    @staticmethod
    def backward(ctx, du):
        sxyzs, nidxs, nlens, nbegs, ncens = ctx.saved_tensors
        du = du.contiguous()
        csum, H = du.shape
        dpa = ops.neighbor_sum_feat_gather(du, ncens, nlens, nbegs)               # sum over each centre's edges
        dpb = ops.neighbor_gather(du, nidxs, nlens, nbegs, use_diff=False)        # sum over edges pointing at j
        P = int(lib.pcs_edge_xyz_wgrad_partials(csum, H))
        if P == 0:
            dwx = torch.zeros((H, 3), dtype=torch.float32, device=du.device)
        else:
            part = torch.empty((P, H, 3), dtype=torch.float32, device=du.device)
            check(lib.pcs_edge_xyz_wgrad(du.data_ptr(), sxyzs.data_ptr(), csum, H, part.data_ptr(), _stream()),
                  "edge_xyz_wgrad")
            dwx = column_sum(part)
        return torch.cat([dpa, dpb], 1), None, dwx, None, None, None, None


def edge_combine(pab, sxyzs, wx, nidxs, nlens, nbegs, ncens):
    """pab = [pa | pb] [pn,2H] (one point-level GEMM), sxyzs [csum,3], wx [H,3] + the CSR graph -> u [csum,H]
    = pa[centre] + pb[neighbour] + sxyz wx^T; differentiable in pab and wx."""
    if not pab.is_cuda:
        raise ValueError("edge_combine: must be a CUDA tensor (this library has no CPU path)")
    return _EdgeCombine.apply(pab, sxyzs, wx, nidxs, nlens, nbegs, ncens)


class _DenseMLP3Pool(torch.autograd.Function):
    """graph_pool(dense_mlp3(u)): the edge MLP followed by the segmented max (a5), differentiated as one
    op so that dy [csum, O] is never materialised and dWo only touches the arg-max rows."""

    @staticmethod
    def forward(ctx, u, w2, w3, wo, nlens, nbegs, ncens):
        u, w2, w3, wo = u.contiguous(), w2.contiguous(), w3.contiguous(), wo.contiguous()
        B, A = w2.shape
        C, O = w3.shape[0], wo.shape[0]
        rows = u.shape[0]
        if u.shape[1] != A + B + C + O or w3.shape[1] != A + B or wo.shape[1] != A + B + C:
            raise ValueError("dense_mlp3_pool: inconsistent layer widths")
        if ncens.shape[0] != rows or nlens.shape != nbegs.shape:
            raise ValueError("dense_mlp3_pool: graph does not match u")
        y = torch.empty((rows, O), dtype=torch.float32, device=u.device)
        check(lib.pcs_dense_mlp3_forward(u.data_ptr(), w2.data_ptr(), w3.data_ptr(), wo.data_ptr(), rows, A, B, C, O,
                                         y.data_ptr(), _stream()), "dense_mlp3_forward")
        pooled, midx = ops.neighbor_max_feat_gather(y, nlens, nbegs)
        ctx.save_for_backward(u, w2, w3, wo, midx, nlens, nbegs, ncens)
        ctx.mark_non_differentiable(midx)
        return pooled, midx

    @staticmethod
    def backward(ctx, dpool, _):
        u, w2, w3, wo, midx, nlens, nbegs, ncens = ctx.saved_tensors
        B, A = w2.shape
        C, O = w3.shape[0], wo.shape[0]
        HC = A + B + C
        rows, pn = u.shape[0], nlens.shape[0]
        dpool = dpool.contiguous()
        du = torch.empty_like(u)
        hcat = torch.empty((rows, HC), dtype=torch.float32, device=u.device)
        Pw = int(lib.pcs_dense_mlp3_wgrad_partials(rows, A, B, C, O))
        wpart = torch.empty((Pw, C * (A + B) + B * A), dtype=torch.float32, device=u.device) if Pw else None
        check(lib.pcs_dense_mlp3_backward_pooled(u.data_ptr(), dpool.data_ptr(), midx.data_ptr(), ncens.data_ptr(),
                                                 nbegs.data_ptr(), w2.data_ptr(), w3.data_ptr(), wo.data_ptr(), rows,
                                                 A, B, C, O, du.data_ptr(), hcat.data_ptr(),
                                                 wpart.data_ptr() if Pw else None, _stream()),
              "dense_mlp3_backward_pooled")
        P = int(lib.pcs_pool_linear_wgrad_partials(pn, O)) if rows else 0
        if P == 0:
            dwo = torch.zeros_like(wo)
        else:
            part = torch.empty((P, O, HC), dtype=torch.float32, device=u.device)
            check(lib.pcs_pool_linear_wgrad(dpool.data_ptr(), midx.data_ptr(), nlens.data_ptr(), nbegs.data_ptr(),
                                            hcat.data_ptr(), pn, rows, O, HC, part.data_ptr(), _stream()),
                  "pool_linear_wgrad")
            dwo = column_sum(part)
        if Pw:  # accumulated on the tensor cores inside the backward kernel
            wsum = column_sum(wpart)
            dw3, dw2 = wsum[:C * (A + B)].view(C, A + B), wsum[C * (A + B):].view(B, A)
        else:
            dw3 = tn_tall(du[:, A + B:A + B + C], hcat[:, C:])        # [C, B+A]
            dw2 = tn_tall(du[:, A:A + B], hcat[:, C + B:])            # [B, A]
        return du, dw2, dw3, dwo, None, None, None


def dense_mlp3_pool(u, w2, w3, wo, nlens, nbegs, ncens):
    """max over each centre's rows of dense_mlp3(u): -> pooled [pn, O] (and the arg-max offsets)."""
    if not u.is_cuda:
        raise ValueError("dense_mlp3_pool: must be a CUDA tensor (this library has no CPU path)")
    return _DenseMLP3Pool.apply(u, w2, w3, wo, nlens, nbegs, ncens)

"""TEST INFRASTRUCTURE, NOT PRODUCT.  CPU restatement of the libPointUtil calls the reference makes either side of
the hot path: grid down-sample, query-vs-source radius search, k-NN, Gaussian label
interpolation, covariance features, block cropping.

PARITY UNPINNED BY THE REFERENCE: libPointUtil is an un-vendored dependency of liuyuan-pal/PointCloudSegmentation
(imported as a prebuilt module, no source or version in the tree), so there is nothing to execute or to read here.
The semantics below are restated from the call sites (s3dis_util.py:33-89, aug_util.py:46-101,285-298,
interpolate.py:135-142, interpolate_semantic3d_dense.py:87-118) and fixed in include/pcseg_ops.h; this module pins the
CUDA path to that statement, bit-exact for indices, 1e-5 for the float outputs (sums here are float64).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import pcs_oracle as orc

_lib = orc._lib
_lib.orc_cross_search_count.restype = C.c_int64
_p, _i64, _f = orc._p, orc._i64, orc._f


def grid_downsample(xyzs, stride):
    """One point per occupied voxel: the lowest index, voxels in ascending (x, y, z) order (s3dis_util.py:41,64)."""
    xyzs = _f(xyzs)
    mn = xyzs.min(axis=0)
    vidx = np.floor((xyzs - mn) / np.float32(stride)).astype(np.int64)  # fp32 subtract and divide, as the kernel
    key = (vidx[:, 0] << 32) + (vidx[:, 1] << 16) + vidx[:, 2]
    order = np.argsort(key, kind="stable")
    ks = key[order]
    head = np.concatenate([[True], ks[1:] != ks[:-1]])
    return order[head].astype(np.int32)


def cross_search(sxyzs, qxyzs, sq_max, sq_min=None):
    """Sources within the radius of every query, ascending index -> (idxs, lens, begs, cens) (s3dis_util.py:66)."""
    s, q = _f(sxyzs), _f(qxyzs)
    sn, qn = s.shape[0], q.shape[0]
    use_min = C.c_int(0 if sq_min is None else 1)
    smin = orc._sq(0.0 if sq_min is None else sq_min)
    lens = np.zeros(qn, np.int32); begs = np.zeros(qn, np.int32)
    en = _lib.orc_cross_search_count(_p(s), _i64(sn), _p(q), _i64(qn), smin, orc._sq(sq_max), use_min, _p(lens), _p(begs))
    idxs = np.zeros(en, np.int32); cens = np.zeros(en, np.int32)
    _lib.orc_cross_search_fill(_p(s), _i64(sn), _p(q), _i64(qn), smin, orc._sq(sq_max), use_min, _p(begs), _p(idxs), _p(cens))
    return idxs, lens, begs, cens


def knn(sxyzs, qxyzs, K):
    """K nearest sources of every query, ascending (squared distance, index) -> (idxs [qn,K], d2 [qn,K]) (interpolate.py:136)."""
    s, q = _f(sxyzs), _f(qxyzs)
    qn = q.shape[0]
    idxs = np.zeros((qn, K), np.int32); d2 = np.zeros((qn, K), np.float32)
    _lib.orc_knn(_p(s), _i64(s.shape[0]), _p(q), _i64(qn), _i64(K), _p(idxs), _p(d2))
    return idxs, d2


def interpolate_probs(sxyzs, qxyzs, sprobs, nidxs, lens, begs, ratio):
    """Gaussian-weighted mean of the neighbours' rows, w = exp(-d^2 ratio) (interpolate.py:140); float64."""
    s, q, pr = np.asarray(sxyzs, np.float64), np.asarray(qxyzs, np.float64), np.asarray(sprobs, np.float64)
    out = np.zeros((q.shape[0], pr.shape[1]))
    for i in range(q.shape[0]):
        n = nidxs[begs[i]:begs[i] + lens[i]]
        if len(n) == 0:
            continue
        w = np.exp(-((s[n] - q[i]) ** 2).sum(1) * float(np.float32(ratio)))
        out[i] = (w[:, None] * pr[n]).sum(0) / w.sum() if w.sum() > 0 else pr[n].mean(0)
    return out


def compute_covars(xyzs, nidxs, lens, begs):
    """3x3 covariance of every neighbourhood about its mean, unit Frobenius norm (s3dis_util.py:72); float64."""
    x = np.asarray(xyzs, np.float64)
    out = np.zeros((len(lens), 9))
    for i in range(len(lens)):
        p = x[nidxs[begs[i]:begs[i] + lens[i]]]
        if len(p) == 0:
            continue
        d = p - p.mean(0)
        c = d.T @ d
        f = np.sqrt((c * c).sum())
        out[i] = (c / f).reshape(-1) if f > 0 else 0.0
    return out


def block_origins(maxx, block_size, stride):
    """aug_util.get_list_without_back_sample (aug_util.py:46-54)."""
    space = float(maxx) - block_size
    return [0.0] if space < 0 else [float(v) for v in np.arange(0, space, stride)] + [space]


def gather_block_points(xyz, begs, block_size):
    """np.nonzero of the block predicate, per block (aug_util.py:74-80) -> list of int32 arrays."""
    xyz = _f(xyz)
    bs = np.float32(block_size)
    out = []
    for bx, by in np.asarray(begs, np.float32).reshape(-1, 2):
        m = (xyz[:, 0] >= bx) & (xyz[:, 0] < bx + bs) & (xyz[:, 1] >= by) & (xyz[:, 1] < by + bs)
        out.append(np.nonzero(m)[0].astype(np.int32))
    return out


def sample_block(points, labels, ds_stride, block_size, block_stride, min_pn, covar_ds_stride=0.03, covar_nn_size=0.1):
    """s3dis_util.sample_block (s3dis_util.py:33-89) without the augmentation flags."""
    points = _f(points)
    xyzs, rgbs = points[:, :3], points[:, 3:]
    mn = xyzs.min(axis=0, keepdims=True)
    c_idx = grid_downsample(xyzs, covar_ds_stride)
    c_xyzs = np.ascontiguousarray(xyzs[c_idx])
    d_idx = grid_downsample(c_xyzs, ds_stride)
    nidxs, nlens, nbgs, _ = cross_search(c_xyzs, c_xyzs[d_idx], covar_nn_size * covar_nn_size)
    covars = compute_covars(c_xyzs, nidxs, nlens, nbgs)
    xyzs_d, rgbs_d, lbls_d = c_xyzs[d_idx], rgbs[c_idx][d_idx], np.asarray(labels)[c_idx][d_idx]
    norm = xyzs_d - mn
    mx = norm.max(axis=0)
    begs = [(x, y) for x in block_origins(mx[0], block_size, block_stride) for y in block_origins(mx[1], block_size, block_stride)]
    blocks = [b for b in gather_block_points(norm, begs, block_size) if len(b) >= min_pn]
    return ([xyzs_d[b] for b in blocks], [rgbs_d[b] for b in blocks], [covars[b] for b in blocks], [lbls_d[b] for b in blocks])

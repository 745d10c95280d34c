"""The libPointUtil calls the reference makes either side of the hot path, on the GPU:
block preparation (s3dis_util.py:33-89, aug_util.py:57-101) and label interpolation (interpolate.py:135-142,
interpolate_semantic3d_dense.py:87-94).  Function names and argument orders are libPointUtil's as the reference
calls them; neighbour lists come back as the CSR the hot path uses (nidxs, lens, bgs[, cens]) instead of Python
lists of arrays, index tensors are int32 CUDA tensors.

libPointUtil is an un-vendored dependency of the reference (no source, no pinned version): the semantics are
restated from the call sites and fixed in include/pcseg_ops.h — parity is pinned by the CPU restatement the tests
carry, not by the reference.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from ._lib import check, lib
from .ops import _f32only, _idx, _ptr, _req, _stream, _workspace

__all__ = ["sortVoxelGPU", "gridDownsampleGPU", "findNeighborRadiusGPU", "findNeighborRadiusCPU", "findNeighborInAnotherCPU",
           "interpolateProbsGPU", "computeCovarsGPU", "gatherBlockPointsGPU", "sample_block", "interpolate",
           "uniform_sample_block"]


def _voxel_index_from_min(xyzs, voxel_len):
    """floor((p - min(p)) / voxel_len) with the IEEE divide of ComputeVoxelIndex; one host sync for the corner."""
    pn = xyzs.shape[0]
    mn = xyzs.min(dim=0).values.tolist()
    out = torch.empty((pn, 3), dtype=torch.int32, device=xyzs.device)
    check(lib.pcs_compute_voxel_index(_ptr(xyzs), pn, 3, float(np.float32(voxel_len)), mn[0], mn[1], mn[2], _ptr(out),
                                      _stream()), "compute_voxel_index")
    return out


def sortVoxelGPU(xyzs, voxel_len):
    """aug_util.py:361: (sidxs, vlens) = the permutation that sorts the points by voxel (ascending (x, y, z) voxel key,
    stable) and the number of points of every occupied voxel, in that order."""
    xyzs = _f32only(xyzs, "xyzs")
    o2p, _, lens, _, _ = ops.compute_permutation_info(_voxel_index_from_min(xyzs, voxel_len))
    return o2p, lens


def gridDownsampleGPU(xyzs, sample_stride, shuffle=False):
    """-> idxs [vn] int32: ONE point per occupied voxel of edge `sample_stride` (s3dis_util.py:41,64).  Here: the point
    with the lowest index in each voxel, voxels in ascending (x, y, z) key order — so the kept points stay spatially
    coherent in index order, which is what the culled searches downstream want.  `shuffle` must be False (the only
    value the reference passes)."""
    _req(not shuffle, "gridDownsampleGPU: shuffle=True is not supported")
    xyzs = _f32only(xyzs, "xyzs")
    _req(xyzs.shape[1] == 3, "xyzs dim 1")
    if xyzs.shape[0] == 0:
        return torch.empty((0,), dtype=torch.int32, device=xyzs.device)
    o2p, _, _, begs, _ = ops.compute_permutation_info(_voxel_index_from_min(xyzs, sample_stride))
    return o2p[begs.long()].contiguous()  # stable sort: the first point of a run has the lowest original index


def findNeighborRadiusGPU(xyzs, radius, leafsize=None):
    """All points within `radius` of every point (aug_util.py:285-298) -> (nidxs, lens, bgs, cens); rows ascending,
    self included (the a18 search)."""
    return ops.search_neighborhood_brute_force(xyzs, radius * radius)


def _cross_search(sxyzs, qxyzs, sq_max, sq_min=None):
    sxyzs, qxyzs = _f32only(sxyzs, "sxyzs"), _f32only(qxyzs, "qxyzs")
    _req(sxyzs.shape[1] == 3 and qxyzs.shape[1] == 3, "xyzs dim 1")
    sn, qn = sxyzs.shape[0], qxyzs.shape[0]
    dev = sxyzs.device
    lens = torch.empty((qn,), dtype=torch.int32, device=dev)
    begs = torch.empty((qn,), dtype=torch.int32, device=dev)
    total = torch.zeros((1,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.pcs_cross_search_workspace_bytes(sn, qn), dev)
    f = lambda v: float(np.float32(v))  # noqa: E731
    use_min = 0 if sq_min is None else 1
    smin = 0.0 if sq_min is None else f(sq_min)
    check(lib.pcs_cross_search_count(_ptr(sxyzs), sn, _ptr(qxyzs), qn, smin, f(sq_max), use_min, _ptr(lens), _ptr(begs),
                                     _ptr(total), _ptr(ws), ws.numel(), _stream()), "cross_search_count")
    en = int(total.item())  # the sizing sync (as SearchNeighborhood.cc:76-83 reads its count back)
    idxs = torch.empty((en,), dtype=torch.int32, device=dev)
    cens = torch.empty((en,), dtype=torch.int32, device=dev)
    check(lib.pcs_cross_search_fill(_ptr(sxyzs), sn, _ptr(qxyzs), qn, smin, f(sq_max), use_min, _ptr(begs), _ptr(idxs),
                                    _ptr(cens), _ptr(ws), ws.numel(), _stream()), "cross_search_fill")
    return idxs, lens, begs, cens


def findNeighborRadiusCPU(xyzs, q_idxs_or_radius, radius=None, leafsize=None):
    """findNeighborRadiusCPU(xyzs, radius[, leafsize]) = findNeighborRadiusGPU; findNeighborRadiusCPU(xyzs, q_idxs,
    radius) (s3dis_util.py:66) = neighbours, in the whole cloud, of the points xyzs[q_idxs] -> CSR over the queries."""
    if not isinstance(q_idxs_or_radius, torch.Tensor):
        return findNeighborRadiusGPU(xyzs, q_idxs_or_radius)
    xyzs = _f32only(xyzs, "xyzs")
    q = _idx(q_idxs_or_radius, "q_idxs")
    return _cross_search(xyzs, xyzs[q.long()].contiguous(), radius * radius)


def findNeighborInAnotherCPU(sxyzs, qxyzs, k_or_radius, return_dists=False):
    """int K: the K nearest sources of every query (interpolate.py:136) -> idxs [qn, K] int32, ascending by (squared
    distance, index) [, squared distances].  float radius: every source within it (interpolate_semantic3d_dense.py:111)
    -> CSR."""
    if isinstance(k_or_radius, float):
        return _cross_search(sxyzs, qxyzs, k_or_radius * k_or_radius)
    K = int(k_or_radius)
    sxyzs, qxyzs = _f32only(sxyzs, "sxyzs"), _f32only(qxyzs, "qxyzs")
    _req(sxyzs.shape[1] == 3 and qxyzs.shape[1] == 3, "xyzs dim 1")
    sn, qn = sxyzs.shape[0], qxyzs.shape[0]
    _req(1 <= K <= 32, "K must be in [1, 32]")
    _req(sn >= K, "fewer sources than K")
    dev = sxyzs.device
    # first radius: a ball that holds ~2K sources at the mean density of the sources' bounding box (one host sync)
    ext = (sxyzs.max(dim=0).values - sxyzs.min(dim=0).values).clamp(min=1e-6).tolist()
    ext.sort()
    dens_vol = ext[0] * ext[1] * ext[2] / sn
    dens_area = ext[1] * ext[2] / sn  # surface-like clouds: the two larger extents
    r0 = float(max(min((2 * K * dens_vol * 3 / (4 * np.pi)) ** (1 / 3), (2 * K * dens_area / np.pi) ** 0.5), 1e-4))
    idxs = torch.empty((qn, K), dtype=torch.int32, device=dev)
    dists = torch.empty((qn, K), dtype=torch.float32, device=dev) if return_dists else None
    rad2 = torch.empty((max(qn, 1),), dtype=torch.float32, device=dev)
    ws = _workspace(lib.pcs_cross_search_workspace_bytes(sn, 0), dev)
    check(lib.pcs_knn_cross(_ptr(sxyzs), sn, _ptr(qxyzs), qn, K, r0, 6, _ptr(idxs), _ptr(dists), _ptr(rad2), _ptr(ws),
                            ws.numel(), _stream()), "knn_cross")
    return (idxs, dists) if return_dists else idxs


def interpolateProbsGPU(sxyzs, qxyzs, sprobs, nidxs, nidxs_lens, nidxs_bgs, ratio):
    """interpolate.py:140 -> qprobs [qn, classes]: Gaussian-weighted mean of the neighbours' rows, w = exp(-d^2 ratio)."""
    sxyzs, qxyzs, sprobs = _f32only(sxyzs, "sxyzs"), _f32only(qxyzs, "qxyzs"), _f32only(sprobs, "sprobs")
    nidxs, lens, bgs = _idx(nidxs, "nidxs"), _idx(nidxs_lens, "nidxs_lens"), _idx(nidxs_bgs, "nidxs_bgs")
    qn = qxyzs.shape[0]
    _req(sprobs.shape[0] == sxyzs.shape[0], "sprobs dim 0")
    _req(lens.shape[0] == qn and bgs.shape[0] == qn, "nidxs_lens dim 0")
    out = torch.empty((qn, sprobs.shape[1]), dtype=torch.float32, device=sxyzs.device)
    check(lib.pcs_interpolate_probs(_ptr(sxyzs), _ptr(qxyzs), _ptr(sprobs), _ptr(nidxs), _ptr(lens), _ptr(bgs), qn,
                                    sprobs.shape[1], float(np.float32(ratio)), _ptr(out), _stream()), "interpolate_probs")
    return out


def computeCovarsGPU(xyzs, nidxs, nidxs_lens, nidxs_bgs):
    """s3dis_util.py:72 -> covars [len(nidxs_lens), 9]: unit-Frobenius covariance of every neighbourhood."""
    xyzs = _f32only(xyzs, "xyzs")
    nidxs, lens, bgs = _idx(nidxs, "nidxs"), _idx(nidxs_lens, "nidxs_lens"), _idx(nidxs_bgs, "nidxs_bgs")
    qn = lens.shape[0]
    _req(bgs.shape[0] == qn, "nidxs_bgs dim 0")
    out = torch.empty((qn, 9), dtype=torch.float32, device=xyzs.device)
    check(lib.pcs_compute_covars(_ptr(xyzs), _ptr(nidxs), _ptr(lens), _ptr(bgs), qn, _ptr(out), _stream()), "compute_covars")
    return out


def gatherBlockPointsGPU(xyz, beg_list, block_size):
    """aug_util.py:99 -> (idxs, lens [nb], bgs [nb]): for every block b the indices (ascending) of the points with
    beg_list[b] <= (x, y) < beg_list[b] + block_size."""
    xyz = _f32only(xyz, "xyz")
    dev = xyz.device
    begs = torch.as_tensor(np.asarray(beg_list, dtype=np.float32).reshape(-1, 2)).to(dev).contiguous()
    nb, pn = begs.shape[0], xyz.shape[0]
    nt = lib.pcs_block_members_tiles(pn)
    counts = torch.zeros((nb, max(nt, 1)), dtype=torch.int32, device=dev)
    check(lib.pcs_block_members_count(_ptr(xyz), pn, _ptr(begs), nb, float(np.float32(block_size)), _ptr(counts), _stream()),
          "block_members_count")
    flat = counts.reshape(-1).to(torch.int64)
    incl = torch.cumsum(flat, 0)
    offs = (incl - flat).contiguous()
    lens = counts.sum(dim=1, dtype=torch.int64)
    total = int(incl[-1].item()) if flat.numel() else 0
    idxs = torch.empty((total,), dtype=torch.int32, device=dev)
    check(lib.pcs_block_members_fill(_ptr(xyz), pn, _ptr(begs), nb, float(np.float32(block_size)), _ptr(offs), _ptr(idxs),
                                     _stream()), "block_members_fill")
    bgs = (torch.cumsum(lens, 0) - lens)
    return idxs, lens.to(torch.int32), bgs.to(torch.int32)


# ---- composites ------------------------------------------------------------------------------------
def _list_without_back_sample(maxx, block_size, stride):
    """aug_util.get_list_without_back_sample (aug_util.py:46-54): origins 0, stride, ... < maxx - block_size, then
    maxx - block_size itself (the last block ends at the cloud's edge); a cloud narrower than a block gets origin 0."""
    space = float(maxx) - block_size
    if space < 0:
        return [0.0]
    return [float(v) for v in np.arange(0, space, stride)] + [space]


def uniform_sample_block(xyz, block_size=3.0, stride=1.5, min_pn=2048):
    """aug_util.uniform_sample_block_gpu (aug_util.py:85-101) on min-normalised points -> list of int32 index tensors
    (blocks with fewer than min_pn points dropped)."""
    mx = xyz.max(dim=0).values.tolist()
    begs = [(x, y) for x in _list_without_back_sample(mx[0], block_size, stride)
            for y in _list_without_back_sample(mx[1], block_size, stride)]
    idxs, lens, bgs = gatherBlockPointsGPU(xyz, begs, block_size)
    lens_h, bgs_h = lens.tolist(), bgs.tolist()
    return [idxs[b:b + n] for b, n in zip(bgs_h, lens_h) if n >= min_pn]


def sample_block(points, labels, ds_stride, block_size, block_stride, min_pn, covar_ds_stride=0.03, covar_nn_size=0.1):
    """s3dis_util.sample_block (s3dis_util.py:33-89) without the augmentation flags: grid down-sample the room twice,
    covariance features of the coarse points from their fine neighbourhoods, crop into blocks.  points [n, 6]
    (xyz + rgb) float32, labels [n] -> lists (one entry per block) of xyzs, rgbs, covars, lbls."""
    xyzs = points[:, :3].contiguous()
    rgbs = points[:, 3:].contiguous()
    mn = xyzs.min(dim=0, keepdim=True).values
    c_idx = gridDownsampleGPU(xyzs, covar_ds_stride)
    c_xyzs = xyzs[c_idx.long()].contiguous()
    d_idx = gridDownsampleGPU(c_xyzs, ds_stride)
    nidxs, nlens, nbgs, _ = findNeighborRadiusCPU(c_xyzs, d_idx, covar_nn_size)
    covars = computeCovarsGPU(c_xyzs, nidxs, nlens, nbgs)
    sel = c_idx.long()[d_idx.long()]
    xyzs_d, rgbs_d, lbls_d = c_xyzs[d_idx.long()], rgbs[sel], labels[sel]
    blocks = uniform_sample_block((xyzs_d - mn).contiguous(), block_size, block_stride, min_pn)
    out = ([], [], [], [])
    for b in blocks:
        bl = b.long()
        for lst, t in zip(out, (xyzs_d, rgbs_d, covars, lbls_d)):
            lst.append(t[bl])
    return out


def interpolate(sxyzs, sprobs, qxyzs, ratio=1.0 / (2 * 0.075 * 0.075), K=6):
    """interpolate.py:135-142: K nearest sampled points of every query, Gaussian-weighted vote -> qprobs [qn, classes]."""
    nidxs = findNeighborInAnotherCPU(sxyzs, qxyzs, K)
    qn = qxyzs.shape[0]
    lens = torch.full((qn,), K, dtype=torch.int32, device=qxyzs.device)
    bgs = torch.arange(qn, dtype=torch.int32, device=qxyzs.device) * K
    return interpolateProbsGPU(sxyzs, qxyzs, sprobs, nidxs.reshape(-1), lens, bgs, ratio)

"""Raw (non-differentiable) operator calls: PyTorch CUDA tensors in, libpcseg_ops.so through ctypes,
PyTorch CUDA tensors out.  One function per TensorFlow custom op of the reference,
same positional argument order, same shape checks and the same error strings as the
`OP_REQUIRES(..., errors::InvalidArgument("..."))` of the op's `.cc` wrapper (cited per function);
a failed check raises ValueError where TF raised InvalidArgumentError.

PyTorch owns every buffer (outputs and workspaces come from its caching allocator); all kernels are
enqueued on the caller's current stream.  There is no CPU path: tensors must live on a CUDA device.
"""
from __future__ import annotations

import numpy as np
import torch

from ._lib import check, lib

__all__ = [
    "neighbor_scatter", "neighbor_gather", "neighbor_sum_feat_gather", "neighbor_sum_feat_scatter",
    "neighbor_max_feat_gather", "neighbor_max_feat_scatter", "location_weight_sum",
    "location_weight_sum_backward", "location_weight_feat_sum", "location_weight_feat_sum_backward",
    "eliminate_center", "compute_voxel_index", "compute_permutation_info",
    "compute_permutation_info_with_class", "permutate_feature", "permutate_feature_backward",
    "compute_diff_xyz", "compute_repermutation_info", "search_neighborhood_brute_force",
    "search_neighborhood_brute_force_range", "search_neighborhood_fixed_brute_force",
    "search_neighborhood_fixed_brute_force_range", "search_neighborhood_multi", "csr_transpose", "csr_transpose_symmetric", "transpose_verified", "csr_cidxs", "exclusive_scan",
    "FEATURE_DTYPES", "set_option",
]


# ---- argument plumbing ---------------------------------------------------------------------------
def _stream() -> int:
    # raw handle of the current stream: torch.cuda.current_stream() builds a Stream object (~5 us per op call)
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def _ptr(t: torch.Tensor | None):
    return None if t is None else t.data_ptr()


def _req(cond: bool, msg: str) -> None:
    if not cond:
        raise ValueError(msg)


# storage types of the feature movers (a1-a10, a15): fp32 (the reference's only instantiation) or bf16
# (16-byte lanes of 8 columns, fp32 accumulation; north_star "bf16x8")
FEATURE_DTYPES = (torch.float32, torch.bfloat16)


def _feat(t: torch.Tensor, name: str, ndim: int = 2, like: torch.Tensor | None = None) -> torch.Tensor:
    _req(isinstance(t, torch.Tensor), f"{name}: expected a torch.Tensor")
    _req(t.is_cuda, f"{name}: must be a CUDA tensor (this library has no CPU path)")
    if like is None:
        _req(t.dtype in FEATURE_DTYPES, f"{name}: expected float32 or bfloat16, got {t.dtype}")
    else:
        _req(t.dtype == like.dtype, f"{name}: expected {like.dtype} like the op's first tensor, got {t.dtype}")
    _req(t.dim() == ndim, f"{name}: expected rank {ndim}, got {t.dim()}")
    return t.contiguous()


def _f32only(t: torch.Tensor, name: str) -> torch.Tensor:
    _req(isinstance(t, torch.Tensor) and t.is_cuda, f"{name}: must be a CUDA tensor (this library has no CPU path)")
    _req(t.dtype == torch.float32, f"{name}: expected float32, got {t.dtype}")
    _req(t.dim() == 2, f"{name}: expected rank 2, got {t.dim()}")
    return t.contiguous()


def _entry(name: str, t: torch.Tensor):
    """The C-ABI entry for this storage type: pcs_<name> (fp32) or pcs_<name>_bf16."""
    return getattr(lib, f"pcs_{name}_bf16" if t.dtype == torch.bfloat16 else f"pcs_{name}")


def _idx(t: torch.Tensor, name: str, ndim: int = 1) -> torch.Tensor:
    _req(isinstance(t, torch.Tensor), f"{name}: expected a torch.Tensor")
    _req(t.is_cuda, f"{name}: must be a CUDA tensor (this library has no CPU path)")
    _req(t.dtype == torch.int32, f"{name}: expected int32, got {t.dtype}")
    _req(t.dim() == ndim, f"{name}: expected rank {ndim}, got {t.dim()}")
    return t.contiguous()


def _workspace(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def _f32(x: float) -> float:
    """A Python double narrowed to fp32, the way a TF `float` attr reaches the kernel."""
    return float(np.float32(x))


def set_option(name: str, value: int) -> None:
    """Tuning knob of the kernel dispatch (csrc/common.cu: thresholds between kernel variants), for A/B runs."""
    check(lib.pcs_set_option(name.encode(), int(value)), f"set_option({name})")


# ---- libTFNeighborOps.so -------------------------------------------------------------------------
def neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False):
    """NeighborScatter (TFNeighborScatter.cc:12-74): [pn,ifn] -> [csum,ifn]."""
    ifeats = _feat(ifeats, "ifeats"); nidxs = _idx(nidxs, "nidxs")
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    pn, ifn = ifeats.shape
    csum = nidxs.shape[0]
    _req(nidxs_lens.shape[0] == pn, "nidxs_lens dim 0")
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    sfeats = torch.empty((csum, ifn), dtype=ifeats.dtype, device=ifeats.device)
    # narrow rows run edge-parallel and need the centre of every edge; wider rows walk per centre
    # (thresholds: csrc/neighbor_ops.cu, measured by tools/ab_rows.py)
    cidxs = csr_cidxs(nidxs_lens, nidxs_bgs, csum) if (use_diff and ifn * ifeats.element_size() < 48) else None
    check(_entry("neighbor_scatter", ifeats)(_ptr(ifeats), _ptr(nidxs), _ptr(nidxs_lens), _ptr(nidxs_bgs), _ptr(cidxs), pn,
                                   ifn, csum, int(bool(use_diff)), _ptr(sfeats), _stream()), "neighbor_scatter")
    return sfeats


def _cache_get(t, attr, key):
    cached = getattr(t, attr, None)
    return cached[1] if cached is not None and cached[0] == key else None


def _cache_put(t, attr, key, value):
    try:
        setattr(t, attr, (key, value))
    except Exception:  # pragma: no cover - exotic tensor subclasses
        pass
    return value


def csr_cidxs(nidxs_lens, nidxs_bgs, csum):
    """cidxs[bg_i+j] = i (the `cens` the searches return), derived once per graph and cached on the
    `nidxs_bgs` tensor: lets NeighborScatter(use_diff) run edge-parallel although its reference
    signature carries no centre-per-edge array."""
    key = (nidxs_bgs.data_ptr(), nidxs_bgs._version, nidxs_lens.data_ptr(), nidxs_lens._version, int(csum))
    hit = _cache_get(nidxs_bgs, "_pcs_cidxs", key)
    if hit is not None:
        return hit
    pn = nidxs_lens.shape[0]
    cidxs = torch.zeros((csum,), dtype=torch.int32, device=nidxs_bgs.device)
    check(lib.pcs_csr_cidxs(_ptr(nidxs_lens), _ptr(nidxs_bgs), pn, csum, _ptr(cidxs), _stream()), "csr_cidxs")
    return _cache_put(nidxs_bgs, "_pcs_cidxs", key, cidxs)


def csr_transpose(nidxs, pn):
    """Edges grouped by neighbour, ascending edge id: (tperm[csum], tbgs[pn+1]).  Cached on
    the `nidxs` tensor object — a radius graph is reused by every layer of a stage, so the sort is paid
    once per graph.  (No reference counterpart: the reference scatters with atomicAdd.)"""
    key = (nidxs.data_ptr(), nidxs._version, int(nidxs.shape[0]), int(pn))
    hit = _cache_get(nidxs, "_pcs_transpose", key)
    if hit is not None:
        # an entry written by csr_transpose_symmetric with a caller-owned flag is trusted only once that flag has
        # been seen to be zero (a non-symmetric graph leaves placeholders in tperm): validate lazily, once
        flag = getattr(nidxs, "_pcs_transpose_flag", None)
        if flag is None:
            return hit
        nidxs._pcs_transpose_flag = None
        if int(flag.item()) == 0:
            return hit
    csum = nidxs.shape[0]
    tperm = torch.empty((csum,), dtype=torch.int32, device=nidxs.device)
    tbgs = torch.empty((pn + 1,), dtype=torch.int32, device=nidxs.device)
    nbytes = lib.pcs_csr_transpose_workspace_bytes(pn, csum)
    ws = _workspace(nbytes, nidxs.device)
    check(lib.pcs_csr_transpose(_ptr(nidxs), pn, csum, _ptr(tperm), _ptr(tbgs), _ptr(ws), ws.numel(),
                                _stream()), "csr_transpose")
    return _cache_put(nidxs, "_pcs_transpose", key, (tperm, tbgs))


def csr_transpose_symmetric(nidxs, nidxs_lens, nidxs_bgs, cidxs, fail_flag=None):
    """csr_transpose for a symmetric CSR with ascending rows (the radius / annulus searches' output): one
    binary search per edge instead of a radix sort.  Fills the same per-tensor cache csr_transpose reads.
    `fail_flag` (int32 CUDA scalar, zeroed) is set by the kernel when the graph is not symmetric; without one
    the flag is read back here (a host sync) and a ValueError raised."""
    nidxs = _idx(nidxs, "nidxs"); nidxs_lens = _idx(nidxs_lens, "nidxs_lens")
    nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs"); cidxs = _idx(cidxs, "cidxs")
    pn, csum = nidxs_lens.shape[0], nidxs.shape[0]
    _req(nidxs_bgs.shape[0] == pn and cidxs.shape[0] == csum, "csr_transpose_symmetric: size mismatch")
    key = (nidxs.data_ptr(), nidxs._version, int(csum), int(pn))
    hit = _cache_get(nidxs, "_pcs_transpose", key)
    if hit is not None:
        return hit
    tperm = torch.empty((csum,), dtype=torch.int32, device=nidxs.device)
    tbgs = torch.empty((pn + 1,), dtype=torch.int32, device=nidxs.device)
    own = fail_flag is None
    if own:
        fail_flag = torch.zeros((), dtype=torch.int32, device=nidxs.device)
    check(lib.pcs_csr_transpose_symmetric(_ptr(nidxs), _ptr(nidxs_lens), _ptr(nidxs_bgs), _ptr(cidxs), pn, csum,
                                          _ptr(tperm), _ptr(tbgs), _ptr(fail_flag), _stream()), "csr_transpose_symmetric")
    if own and int(fail_flag.item()):
        raise ValueError("csr_transpose_symmetric: the CSR is not symmetric (an edge has no mirror edge)")
    try:  # caller-owned flag: not checked yet — csr_transpose validates it on first use (transpose_verified skips that)
        nidxs._pcs_transpose_flag = None if own else fail_flag
    except Exception:  # pragma: no cover
        pass
    return _cache_put(nidxs, "_pcs_transpose", key, (tperm, tbgs))


def transpose_verified(nidxs):
    """The caller has read csr_transpose_symmetric's fail flag of this graph and found it zero."""
    try:
        nidxs._pcs_transpose_flag = None
    except Exception:  # pragma: no cover
        pass


def neighbor_gather(sfeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False):
    """NeighborGather (TFNeighborGather.cc:12-73): [csum,ifn] -> [pn,ifn]; deterministic."""
    sfeats = _feat(sfeats, "sfeats"); nidxs = _idx(nidxs, "nidxs")
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    pn = nidxs_lens.shape[0]
    csum = nidxs.shape[0]
    ifn = sfeats.shape[1]
    _req(sfeats.shape[0] == csum, "sfeats dim 0")
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    tperm, tbgs = csr_transpose(nidxs, pn)
    ifeats = torch.empty((pn, ifn), dtype=sfeats.dtype, device=sfeats.device)
    check(_entry("neighbor_gather", sfeats)(_ptr(sfeats), _ptr(nidxs_lens), _ptr(nidxs_bgs), _ptr(tperm), _ptr(tbgs),
                                  pn, ifn, csum, int(bool(use_diff)), _ptr(ifeats), _stream()),
          "neighbor_gather")
    return ifeats


def neighbor_sum_feat_gather(ifeats, cidxs, nidxs_lens, nidxs_bgs):
    """NeighborSumFeatGather (TFNeighborSumFeatGather.cc:12-70): [csum,fd] -> [pn,fd]."""
    ifeats = _feat(ifeats, "ifeats"); cidxs = _idx(cidxs, "cidxs")
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    csum, fd = ifeats.shape
    pn = nidxs_lens.shape[0]
    _req(cidxs.shape[0] == csum, "cidxs dim 0")
    _req(nidxs_bgs.shape[0] == pn, "nidixs_bgs dim 0")
    out = torch.empty((pn, fd), dtype=ifeats.dtype, device=ifeats.device)
    check(_entry("neighbor_sum_feat_gather", ifeats)(_ptr(ifeats), _ptr(nidxs_lens), _ptr(nidxs_bgs), pn, fd, csum,
                                           _ptr(out), _stream()), "neighbor_sum_feat_gather")
    return out


def neighbor_sum_feat_scatter(ifeats, cidxs, nidxs_lens, nidxs_bgs):
    """NeighborSumFeatScatter (TFNeighborSumFeatScatter.cc:12-61): [pn,fd] -> [csum,fd]."""
    ifeats = _feat(ifeats, "ifeats"); cidxs = _idx(cidxs, "cidxs")
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    pn, fd = ifeats.shape
    csum = cidxs.shape[0]
    out = torch.empty((csum, fd), dtype=ifeats.dtype, device=ifeats.device)
    _req(nidxs_lens.shape[0] == pn, "nidxs_lens dim 0")
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    check(_entry("neighbor_sum_feat_scatter", ifeats)(_ptr(ifeats), _ptr(cidxs), _ptr(nidxs_lens), _ptr(nidxs_bgs), pn, fd,
                                            csum, _ptr(out), _stream()), "neighbor_sum_feat_scatter")
    return out


def neighbor_max_feat_gather(ifeats, vlens, vlens_bgs):
    """NeighborMaxFeatGather (TFNeighborMaxFeatGather.cc:12-69): [pn1,fd] -> ([pn2,fd] f32, [pn2,fd] i32)."""
    ifeats = _feat(ifeats, "ifeats"); vlens = _idx(vlens, "vlens"); vlens_bgs = _idx(vlens_bgs, "vlens_bgs")
    pn1, fd = ifeats.shape
    pn2 = vlens.shape[0]
    _req(vlens_bgs.shape[0] == pn2, "vlens_bgs dim 0")
    gfeats = torch.empty((pn2, fd), dtype=ifeats.dtype, device=ifeats.device)
    max_idxs = torch.empty((pn2, fd), dtype=torch.int32, device=ifeats.device)
    check(_entry("neighbor_max_feat_gather", ifeats)(_ptr(ifeats), _ptr(vlens), _ptr(vlens_bgs), pn1, pn2, fd,
                                           _ptr(gfeats), _ptr(max_idxs), _stream()), "neighbor_max_feat_gather")
    return gfeats, max_idxs


def neighbor_max_feat_scatter(gfeats, ifeats, max_idxs, vlens_bgs, vlens):
    """NeighborMaxFeatScatter (TFNeighborMaxFeatScatter.cc:12-80) — note the (bgs, lens) order.
    `ifeats` only supplies pn1 (a tensor, or an int for callers that do not keep the features)."""
    gfeats = _feat(gfeats, "gfeats"); max_idxs = _idx(max_idxs, "max_idxs", 2)
    vlens_bgs = _idx(vlens_bgs, "vlens_bgs"); vlens = _idx(vlens, "vlens")
    pn2, fd = gfeats.shape
    if isinstance(ifeats, int):
        pn1 = ifeats
    else:
        _req(ifeats.dim() == 2 and ifeats.shape[1] == fd, "ifeats dim 1")
        pn1 = ifeats.shape[0]
    _req(max_idxs.shape[0] == pn2, "max_idxs dim 0")
    _req(max_idxs.shape[1] == fd, "max_idxs dim 1")
    _req(vlens_bgs.shape[0] == pn2, "vlens_bgs dim 0")
    _req(vlens.shape[0] == pn2, "vlens dim 0")
    sfeats = torch.empty((pn1, fd), dtype=gfeats.dtype, device=gfeats.device)
    check(_entry("neighbor_max_feat_scatter", gfeats)(_ptr(gfeats), _ptr(max_idxs), _ptr(vlens_bgs), _ptr(vlens), pn1,
                                            pn2, fd, _ptr(sfeats), _stream()), "neighbor_max_feat_scatter")
    return sfeats


def location_weight_sum(lw, nidxs_lens, nidxs_bgs):
    """LocationWeightSum (TFLocationWeightSumForward.cc:12-65): [csum,m] -> [pn,m]."""
    lw = _feat(lw, "lw"); nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    csum, m = lw.shape
    pn = nidxs_lens.shape[0]
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    out = torch.empty((pn, m), dtype=lw.dtype, device=lw.device)
    check(_entry("location_weight_sum", lw)(_ptr(lw), _ptr(nidxs_lens), _ptr(nidxs_bgs), pn, m, csum, _ptr(out),
                                      _stream()), "location_weight_sum")
    return out


def location_weight_sum_backward(lw, dlw_sum, nidxs_lens, nidxs_bgs):
    """LocationWeightSumBackward (TFLocationWeightSumBackward.cc:12-68); `lw` only supplies the shape."""
    dlw_sum = _feat(dlw_sum, "dlw_sum"); nidxs_lens = _idx(nidxs_lens, "nidxs_lens")
    nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    _req(lw.dim() == 2, "lw: expected rank 2")
    csum, m = lw.shape
    pn = nidxs_lens.shape[0]
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    _req(dlw_sum.shape[0] == pn, "dlw_sum dim 0")
    _req(dlw_sum.shape[1] == m, "dlw_sum dim 1")
    dlw = torch.empty((csum, m), dtype=dlw_sum.dtype, device=dlw_sum.device)
    check(_entry("location_weight_sum_backward", dlw_sum)(_ptr(dlw_sum), _ptr(nidxs_lens), _ptr(nidxs_bgs), pn, m, csum,
                                               _ptr(dlw), _stream()), "location_weight_sum_backward")
    return dlw


def location_weight_feat_sum(lw, feats, nidxs_lens, nidxs_bgs):
    """LocationWeightFeatSum (TFLocationWeightFeatSumForward.cc:12-72): [csum,m],[csum,m,ofn] -> [pn,m,ofn]."""
    lw = _feat(lw, "lw"); feats = _feat(feats, "feats", 3, like=lw)
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    csum, m, ofn = feats.shape
    pn = nidxs_lens.shape[0]
    _req(lw.shape[0] == csum, "lw dim 0")
    _req(lw.shape[1] == m, "lw dim 1")
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    out = torch.empty((pn, m, ofn), dtype=feats.dtype, device=feats.device)
    check(_entry("location_weight_feat_sum", lw)(_ptr(lw), _ptr(feats), _ptr(nidxs_lens), _ptr(nidxs_bgs), pn, m, ofn,
                                           csum, _ptr(out), _stream()), "location_weight_feat_sum")
    return out


def location_weight_feat_sum_backward(lw, feats, dtfeats_sum, nidxs_lens, nidxs_bgs):
    """LocationWeightFeatSumBackward (TFLocationWeightFeatSumBackward.cc:12-81) -> (dlw, dfeats).
    Edges outside every segment: dlw = 0 (the reference's memset, TFNeighborKernel.cu:409); the matching rows of
    dfeats are NOT written — upstream's memset of ditfeats is commented out too — so they are uninitialised."""
    lw = _feat(lw, "lw"); feats = _feat(feats, "feats", 3, like=lw); dtfeats_sum = _feat(dtfeats_sum, "dtfeats_sum", 3, like=lw)
    nidxs_lens = _idx(nidxs_lens, "nidxs_lens"); nidxs_bgs = _idx(nidxs_bgs, "nidxs_bgs")
    csum, m, ofn = feats.shape
    pn = nidxs_lens.shape[0]
    _req(lw.shape[0] == csum, "lw dim 0")
    _req(lw.shape[1] == m, "lw dim 1")
    _req(nidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    _req(dtfeats_sum.shape[0] == pn, "dtfeats_sum dim 0")
    _req(dtfeats_sum.shape[1] == m, "dtfeats_sum dim 0")  # sic: the reference reuses the string
    _req(dtfeats_sum.shape[2] == ofn, "dtfeats_sum dim 0")
    dlw = torch.zeros((csum, m), dtype=feats.dtype, device=feats.device)
    dfeats = torch.empty((csum, m, ofn), dtype=feats.dtype, device=feats.device)
    check(_entry("location_weight_feat_sum_backward", lw)(_ptr(lw), _ptr(feats), _ptr(dtfeats_sum), _ptr(nidxs_lens),
                                                    _ptr(nidxs_bgs), pn, m, ofn, csum, _ptr(dlw), _ptr(dfeats),
                                                    _stream()), "location_weight_feat_sum_backward")
    return dlw, dfeats


def eliminate_center(inidxs, inidxs_lens, inidxs_bgs):
    """EliminateCenter (TFNeighborEliminateCenter.cc:13-82) -> (onidxs, onidxs_lens, onidxs_bgs, ocidxs).
    Precondition as upstream: every row contains its own centre exactly once."""
    inidxs = _idx(inidxs, "inidxs"); inidxs_lens = _idx(inidxs_lens, "inidxs_lens")
    inidxs_bgs = _idx(inidxs_bgs, "inidxs_bgs")
    csum = inidxs.shape[0]
    pn = inidxs_lens.shape[0]
    _req(inidxs_bgs.shape[0] == pn, "nidxs_bgs dim 0")
    _req(csum >= pn, "eliminate_center: csum < pn (rows must contain their own centre)")
    dev = inidxs.device
    onidxs = torch.empty((csum - pn,), dtype=torch.int32, device=dev)
    ocidxs = torch.empty((csum - pn,), dtype=torch.int32, device=dev)
    olens = torch.empty((pn,), dtype=torch.int32, device=dev)
    obgs = torch.empty((pn,), dtype=torch.int32, device=dev)
    check(lib.pcs_eliminate_center(_ptr(inidxs), _ptr(inidxs_lens), _ptr(inidxs_bgs), pn, csum, _ptr(onidxs),
                                   _ptr(olens), _ptr(obgs), _ptr(ocidxs), _stream()), "eliminate_center")
    return onidxs, olens, obgs, ocidxs


# ---- PoolingOps.so -------------------------------------------------------------------------------
def compute_voxel_index(pts, voxel_len, block_size):
    """ComputeVoxelIndex (PoolingIndex/ComputeVoxelIdx.cc:27-71): [pn,3] f32 -> [pn,3] i32.
    The lower corner is formed in fp32 exactly as the op does (.cc:66-67)."""
    pts = _f32only(pts, "pts")
    _req(pts.shape[1] == 3, "pts dim 1")
    pn = pts.shape[0]
    f = np.float32
    eps = f(1e-3)
    bs = f(block_size)
    mn_xy = -bs / f(2.0) - eps
    mn_z = f(0.0) - eps
    out = torch.empty((pn, 3), dtype=torch.int32, device=pts.device)
    check(lib.pcs_compute_voxel_index(_ptr(pts), pn, 3, _f32(voxel_len), float(mn_xy), float(mn_xy), float(mn_z),
                                      _ptr(out), _stream()), "compute_voxel_index")
    return out


def _permutation_info(voxel_idxs, classes, block_ids):
    voxel_idxs = _idx(voxel_idxs, "voxel_idxs", 2)
    _req(voxel_idxs.shape[1] == 3, "voxel_idxs dim 1")
    pn = voxel_idxs.shape[0]
    dev = voxel_idxs.device
    if classes is not None:
        classes = _idx(classes, "classes")
        _req(classes.shape[0] == pn, "classes dim 0")
    if block_ids is not None:
        block_ids = _idx(block_ids, "block_ids")
        _req(block_ids.shape[0] == pn, "block_ids dim 0")
    o2p = torch.empty((pn,), dtype=torch.int32, device=dev)
    p2o = torch.empty((pn,), dtype=torch.int32, device=dev)
    lens = torch.empty((pn,), dtype=torch.int32, device=dev)
    begs = torch.empty((pn,), dtype=torch.int32, device=dev)
    cens = torch.empty((pn,), dtype=torch.int32, device=dev)
    vn_dev = torch.zeros((2,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.pcs_compute_permutation_info_workspace_bytes(pn), dev)
    args = (_ptr(voxel_idxs), _ptr(classes), _ptr(block_ids), pn, _ptr(o2p), _ptr(p2o), _ptr(lens), _ptr(begs),
            _ptr(cens), _ptr(vn_dev), _ptr(ws), ws.numel(), _stream())
    check(lib.pcs_compute_permutation_info(*args), "compute_permutation_info")
    vn, flags = vn_dev.tolist()  # the one D2H sync: vn sizes the outputs (ComputePermutationInfo.cu:134)
    if flags & 2:
        raise ValueError("compute_permutation_info: voxel coordinate outside [0,65535] with block_ids")
    if flags & 1:  # a key needs more than 48 bits: sort the full 64-bit key like thrust
        check(lib.pcs_compute_permutation_info64(*args), "compute_permutation_info64")
        vn, flags = vn_dev.tolist()
    return o2p, p2o, lens[:vn].contiguous(), begs[:vn].contiguous(), cens


def compute_permutation_info(voxel_idxs, block_ids=None):
    """ComputePermutationInfo, GPU kernel (PoolingIndex/ComputePermutationInfo.cc:121-163)
    -> (o2p[pn], p2o[pn], lens[vn], begs[vn], cens[pn]).  `block_ids` (extension): per-point block id
    for block-diagonal batches, sorted as the most significant key field."""
    return _permutation_info(voxel_idxs, None, block_ids)


def compute_permutation_info_with_class(voxel_idxs, classes):
    """ComputePermutationInfoWithClass (PoolingIndex/ComputePermutationInfo.cc:166-252)."""
    return _permutation_info(voxel_idxs, classes, None)


def permutate_feature(feats, idxs, back_idxs=None):
    """PermutateFeature (PoolingIndex/PermutateFeature.cc:72-121): out[i,:] = feats[idxs[i],:];
    T in {float32, int32}."""
    _req(isinstance(feats, torch.Tensor) and feats.is_cuda, "feats: must be a CUDA tensor")
    _req(feats.dtype in (torch.float32, torch.int32, torch.bfloat16),
         f"feats: expected float32, int32 or bfloat16, got {feats.dtype}")
    _req(feats.dim() == 2, "feats: expected rank 2")
    feats = feats.contiguous()
    idxs = _idx(idxs, "idxs")
    pn, ps = feats.shape
    _req(idxs.shape[0] == pn, "idxs dim 0")
    if back_idxs is not None:
        _req(back_idxs.shape[0] == pn, "back_idxs dim 0")
    out = torch.empty_like(feats)
    check(_entry("permutate_feature", feats)(_ptr(feats), _ptr(idxs), pn, ps, _ptr(out), _stream()), "permutate_feature")
    return out


def permutate_feature_backward(dfeats, back_idxs):
    """PermutateFeatureBackward (PoolingIndex/PermutateFeature.cc:23-69): same kernel, inverse map."""
    return permutate_feature(dfeats, back_idxs)


def compute_diff_xyz(xyzs, cxyzs, cidxs):
    """ComputeDiffXyz (PoolingIndex/ComputeDiffXYZ.cc:21-71): xyzs[i] - cxyzs[cidxs[i]]."""
    xyzs = _f32only(xyzs, "xyzs"); cxyzs = _f32only(cxyzs, "cxyzs"); cidxs = _idx(cidxs, "cidxs")
    _req(xyzs.shape[1] == 3, "xyzs dim 1")
    _req(cxyzs.shape[1] == 3, "cxyzs dim 1")
    pn1, pn2 = xyzs.shape[0], cxyzs.shape[0]
    _req(cidxs.shape[0] == pn1, "cidxs dim 0")
    out = torch.empty_like(xyzs)
    check(lib.pcs_compute_diff_xyz(_ptr(xyzs), _ptr(cxyzs), _ptr(cidxs), pn1, pn2, _ptr(out), _stream()),
          "compute_diff_xyz")
    return out


def compute_repermutation_info(o2p_idxs2, lens, begs, cens):
    """ComputeRepermutationInfo (PoolingIndex/ComputeRepermutationInfo.cc:20-119)
    -> (reper_o2p1[pn1], reper_p2o1[pn1], reper_lens[pn2], reper_begs[pn2], reper_cens[pn1])."""
    o2p_idxs2 = _idx(o2p_idxs2, "o2p_idxs2"); lens = _idx(lens, "lens"); begs = _idx(begs, "begs")
    cens = _idx(cens, "cens")
    pn2, pn1 = o2p_idxs2.shape[0], cens.shape[0]
    _req(lens.shape[0] == pn2, "lens dim 0")
    _req(begs.shape[0] == pn2, "begs dim 0")
    dev = cens.device
    ro2p = torch.empty((pn1,), dtype=torch.int32, device=dev)
    rp2o = torch.empty((pn1,), dtype=torch.int32, device=dev)
    rlens = torch.empty((pn2,), dtype=torch.int32, device=dev)
    rbegs = torch.empty((pn2,), dtype=torch.int32, device=dev)
    rcens = torch.empty((pn1,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.pcs_compute_repermutation_info_workspace_bytes(pn2), dev)
    check(lib.pcs_compute_repermutation_info(_ptr(o2p_idxs2), _ptr(lens), _ptr(begs), pn1, pn2, _ptr(ro2p),
                                             _ptr(rp2o), _ptr(rlens), _ptr(rbegs), _ptr(rcens), _ptr(ws),
                                             ws.numel(), _stream()), "compute_repermutation_info")
    return ro2p, rp2o, rlens, rbegs, rcens


def _blocks(block_begs, pn):
    if block_begs is None:
        return None, 0, 0
    block_begs = _idx(block_begs, "block_begs")
    nb = block_begs.shape[0] - 1
    _req(nb >= 1, "block_begs: need at least one block")
    return block_begs, nb, None


def _search(xyzs, sq_min, sq_max, use_min, block_begs, max_block_pn):
    xyzs = _f32only(xyzs, "xyzs")
    _req(xyzs.shape[1] == 3, "xyzs dim 1")
    pn = xyzs.shape[0]
    dev = xyzs.device
    block_begs, nb, _ = _blocks(block_begs, pn)
    if block_begs is not None and max_block_pn is None:
        max_block_pn = int((block_begs[1:] - block_begs[:-1]).max().item())
    lens = torch.empty((pn,), dtype=torch.int32, device=dev)
    begs = torch.empty((pn,), dtype=torch.int32, device=dev)
    total = torch.zeros((1,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.pcs_search_workspace_bytes(pn), dev)
    mb = int(max_block_pn or 0)
    check(lib.pcs_search_neighborhood_count(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, sq_min, sq_max, use_min,
                                            _ptr(lens), _ptr(begs), _ptr(total), _ptr(ws), ws.numel(), _stream()),
          "search_neighborhood_count")
    en = int(total.item())  # D2H sync sizing the outputs, as SearchNeighborhood.cu:148-149
    idxs = torch.empty((en,), dtype=torch.int32, device=dev)
    cens = torch.empty((en,), dtype=torch.int32, device=dev)
    check(lib.pcs_search_neighborhood_fill(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, sq_min, sq_max, use_min,
                                           _ptr(begs), _ptr(idxs), _ptr(cens), _ptr(ws), ws.numel(), _stream()),
          "search_neighborhood_fill")
    return idxs, lens, begs, cens


def search_neighborhood_brute_force(xyzs, squared_nn_size, block_begs=None, max_block_pn=None):
    """SearchNeighborhoodBruteForce (PoolingIndex/SearchNeighborhood.cc:30-87) -> (idxs, lens, begs, cens).
    `block_begs` (extension): int32[nb+1] point offsets of independent blocks."""
    return _search(xyzs, 0.0, _f32(squared_nn_size), 0, block_begs, max_block_pn)


def search_neighborhood_brute_force_range(xyzs, squared_min_nn_size, squared_max_nn_size, block_begs=None,
                                          max_block_pn=None):
    """SearchNeighborhoodBruteForceRange (PoolingIndex/SearchNeighborhood.cc:109-169)."""
    return _search(xyzs, _f32(squared_min_nn_size), _f32(squared_max_nn_size), 1, block_begs, max_block_pn)


SEARCH_PAD_K = 64  # padded hit-list length of the one-traversal search (0: count and fill both traverse); rows longer than
#                    this are filled by a second traversal of the warps that own them, so any value gives the same output


def search_neighborhood_multi(xyzs, specs, block_begs=None, max_block_pn=None, pad_k=None):
    """Several graphs over the same points from one traversal (extension, csrc/search_multi.cu).  specs = list of
    (squared_min_nn_size or None, squared_max_nn_size); -> list of (idxs, lens, begs, cens), entry g identical to
    search_neighborhood_brute_force[_range] with specs[g].  One count launch, one sizing read for all graphs, one fill
    launch (up to four graphs per launch; longer lists are split).  With pad_k > 0 (default SEARCH_PAD_K) the count
    launch also records each centre's first pad_k hits and the fill step is a row copy."""
    import ctypes as C

    xyzs = _f32only(xyzs, "xyzs")
    _req(xyzs.shape[1] == 3, "xyzs dim 1")
    pn = xyzs.shape[0]
    dev = xyzs.device
    block_begs, nb, _ = _blocks(block_begs, pn)
    if block_begs is not None and max_block_pn is None:
        max_block_pn = int((block_begs[1:] - block_begs[:-1]).max().item())
    mb = int(max_block_pn or 0)
    out = []
    for g0 in range(0, len(specs), 4):
        part = specs[g0:g0 + 4]
        ng = len(part)
        smin = np.array([0.0 if a is None else a for a, _ in part], dtype=np.float32)
        smax = np.array([b for _, b in part], dtype=np.float32)
        umin = np.array([0 if a is None else 1 for a, _ in part], dtype=np.int32)
        lens = torch.empty((ng, pn), dtype=torch.int32, device=dev)
        begs = torch.empty((ng, pn), dtype=torch.int32, device=dev)
        totals = torch.zeros((ng,), dtype=torch.int32, device=dev)
        ws = _workspace(lib.pcs_search_workspace_bytes(pn), dev)
        K = int(SEARCH_PAD_K if pad_k is None else pad_k)
        if K > 0 and pn > 0:
            pad = torch.empty((ng, pn, K), dtype=torch.int32, device=dev)
            check(lib.pcs_search_multi_count_pad(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, ng, smin.ctypes.data, smax.ctypes.data,
                                                 umin.ctypes.data, _ptr(lens), _ptr(begs), _ptr(totals), _ptr(pad), K, _ptr(ws),
                                                 ws.numel(), _stream()), "search_multi_count_pad")
        else:
            check(lib.pcs_search_multi_count(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, ng, smin.ctypes.data, smax.ctypes.data,
                                             umin.ctypes.data, _ptr(lens), _ptr(begs), _ptr(totals), _ptr(ws), ws.numel(),
                                             _stream()), "search_multi_count")
        ens = totals.tolist()  # ONE D2H sync sizes every graph of the call
        idxs = [torch.empty((en,), dtype=torch.int32, device=dev) for en in ens]
        cens = [torch.empty((en,), dtype=torch.int32, device=dev) for en in ens]
        pi = (C.c_void_p * ng)(*[t.data_ptr() for t in idxs])
        pc = (C.c_void_p * ng)(*[t.data_ptr() for t in cens])
        if K > 0 and pn > 0:
            check(lib.pcs_search_multi_fill_pad(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, ng, smin.ctypes.data, smax.ctypes.data,
                                                umin.ctypes.data, _ptr(lens), _ptr(begs), _ptr(pad), K, C.addressof(pi),
                                                C.addressof(pc), _ptr(ws), ws.numel(), _stream()), "search_multi_fill_pad")
        else:
            check(lib.pcs_search_multi_fill(_ptr(xyzs), pn, _ptr(block_begs), nb, mb, ng, smin.ctypes.data, smax.ctypes.data,
                                            umin.ctypes.data, _ptr(begs), C.addressof(pi), C.addressof(pc), _ptr(ws), ws.numel(),
                                            _stream()), "search_multi_fill")
        out += [(idxs[g], lens[g], begs[g], cens[g]) for g in range(ng)]
    return out


def _search_fixed(xyzs, sq_min, sq_max, use_min, fixed_size, block_begs, max_block_pn):
    xyzs = _f32only(xyzs, "xyzs")
    _req(xyzs.shape[1] == 3, "xyzs dim 1")
    pn = xyzs.shape[0]
    dev = xyzs.device
    K = int(fixed_size)
    _req(K >= 0, "fixed_size must be non-negative")
    block_begs, nb, _ = _blocks(block_begs, pn)
    if block_begs is not None and max_block_pn is None:
        max_block_pn = int((block_begs[1:] - block_begs[:-1]).max().item())
    idxs = torch.empty((pn * K,), dtype=torch.int32, device=dev)
    cens = torch.empty((pn * K,), dtype=torch.int32, device=dev)
    lens = torch.empty((pn,), dtype=torch.int32, device=dev)
    begs = torch.empty((pn,), dtype=torch.int32, device=dev)
    ws = _workspace(lib.pcs_search_workspace_bytes(pn), dev)
    check(lib.pcs_search_neighborhood_fixed(_ptr(xyzs), pn, _ptr(block_begs), nb, int(max_block_pn or 0), sq_min,
                                            sq_max, use_min, K, _ptr(idxs), _ptr(lens), _ptr(begs), _ptr(cens),
                                            _ptr(ws), ws.numel(), _stream()), "search_neighborhood_fixed")
    return idxs, lens, begs, cens


def search_neighborhood_fixed_brute_force(xyzs, squared_nn_size, fixed_size, block_begs=None, max_block_pn=None):
    """SearchNeighborhoodFixedBruteForce (PoolingIndex/SearchNeighborhoodFixed.cc:33-101)."""
    return _search_fixed(xyzs, 0.0, _f32(squared_nn_size), 0, fixed_size, block_begs, max_block_pn)


def search_neighborhood_fixed_brute_force_range(xyzs, squared_min_nn_size, squared_max_nn_size, fixed_size,
                                                block_begs=None, max_block_pn=None):
    """SearchNeighborhoodFixedBruteForceRange (PoolingIndex/SearchNeighborhoodFixed.cc:104-166)."""
    return _search_fixed(xyzs, _f32(squared_min_nn_size), _f32(squared_max_nn_size), 1, fixed_size, block_begs,
                         max_block_pn)


def exclusive_scan(x):
    """int32 exclusive prefix sum (thrust::exclusive_scan call sites, SearchNeighborhood.cu:143)."""
    x = _idx(x, "x")
    out = torch.empty_like(x)
    ws = _workspace(lib.pcs_exclusive_scan_workspace_bytes(x.numel()), x.device)
    check(lib.pcs_exclusive_scan_i32(_ptr(x), x.numel(), _ptr(out), None, _ptr(ws), ws.numel(), _stream()),
          "exclusive_scan")
    return out

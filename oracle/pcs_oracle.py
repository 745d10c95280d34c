"""TEST INFRASTRUCTURE, NOT PRODUCT.  numpy-facing ctypes wrapper of oracle/_ref/libpcs_oracle.so (the
CPU restatement in pcs_oracle.c / pcs_oracle_index.c, each function citing the reference file:line it
follows).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; pointcloudsegmentation_b200/ never does.

Parity pinning: tests/test_oracle_golden.py checks these functions against tests/golden/*.npz, which
tests/golden/make_golden.py produced by EXECUTING the reference's own numpy test oracles
(tf_ops/test/*.py) in the authoring container; on the GPU box tests/test_libref_parity.py checks
them (and the product) against oracle/_ref/libref.so = the reference's unmodified .cu files.
Ops with no reference test (location_weight_*, exact search order, voxel index values, repermutation
values) are pinned by libref only.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
SO = _DIR / "_ref" / "libpcs_oracle.so"


def build(quiet: bool = True) -> Path:
    """gcc the restatement (and, when /root/reference exists, libref.so) via oracle/Makefile."""
    r = subprocess.run(["make", "-C", str(_DIR)], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if not quiet:
        print(r.stdout)
    return SO


def _load():
    if not SO.exists():
        r = subprocess.run(["make", "-C", str(_DIR), "_ref/libpcs_oracle.so"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    return C.CDLL(str(SO))


_lib = _load()
_lib.orc_num_threads.restype = C.c_int
_lib.orc_permutation_info.restype = C.c_int64
_lib.orc_search_count.restype = C.c_int64

_i64 = C.c_int64


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def num_threads() -> int:
    return int(_lib.orc_num_threads())


def set_num_threads(n: int) -> None:
    _lib.orc_set_num_threads(C.c_int(int(n)))


# ---- neighbour ops ---------------------------------------------------------------------------------
def neighbor_scatter(ifeats, nidxs, lens, bgs, use_diff=False):
    ifeats, nidxs, lens, bgs = _f(ifeats), _i(nidxs), _i(lens), _i(bgs)
    pn, c = ifeats.shape
    out = np.zeros((nidxs.shape[0], c), np.float32)
    _lib.orc_neighbor_scatter(_p(ifeats), _p(nidxs), _p(lens), _p(bgs), _i64(pn), _i64(c), C.c_int(int(use_diff)), _p(out))
    return out


def neighbor_gather(sfeats, nidxs, lens, bgs, use_diff=False):
    """float64 result (sum accumulated in double)."""
    sfeats, nidxs, lens, bgs = _f(sfeats), _i(nidxs), _i(lens), _i(bgs)
    pn, c = lens.shape[0], sfeats.shape[1]
    out = np.zeros((pn, c), np.float64)
    _lib.orc_neighbor_gather(_p(sfeats), _p(nidxs), _p(lens), _p(bgs), _i64(pn), _i64(c), C.c_int(int(use_diff)), _p(out))
    return out


def segment_sum(feats, lens, bgs):
    """a3 NeighborSumFeatGather / a7 LocationWeightSum; float64 result."""
    feats, lens, bgs = _f(feats), _i(lens), _i(bgs)
    pn, c = lens.shape[0], feats.shape[1]
    out = np.zeros((pn, c), np.float64)
    _lib.orc_segment_sum(_p(feats), _p(lens), _p(bgs), _i64(pn), _i64(c), _p(out))
    return out


def segment_sum_f32(feats, lens, bgs):
    """a3 in the reference's own arithmetic: sequential fp32 `+=` in edge order (TFNeighborKernelNew.cu:19-25)."""
    feats, lens, bgs = _f(feats), _i(lens), _i(bgs)
    pn, c = lens.shape[0], feats.shape[1]
    out = np.zeros((pn, c), np.float32)
    _lib.orc_segment_sum_f32(_p(feats), _p(lens), _p(bgs), _i64(pn), _i64(c), _p(out))
    return out


def row_gather(feats, idx):
    """a4 NeighborSumFeatScatter (idx = cidxs) / a15 PermutateFeature; any 4-byte dtype."""
    feats = np.ascontiguousarray(feats)
    assert feats.dtype.itemsize == 4
    idx = _i(idx)
    out = np.zeros((idx.shape[0], feats.shape[1]), feats.dtype)
    _lib.orc_row_gather(_p(feats), _p(idx), _i64(idx.shape[0]), _i64(feats.shape[1]), _p(out))
    return out


def max_feat_gather(feats, lens, bgs):
    feats, lens, bgs = _f(feats), _i(lens), _i(bgs)
    pn2, c = lens.shape[0], feats.shape[1]
    g = np.zeros((pn2, c), np.float32)
    mi = np.zeros((pn2, c), np.int32)
    _lib.orc_max_feat_gather(_p(feats), _p(lens), _p(bgs), _i64(pn2), _i64(c), _p(g), _p(mi))
    return g, mi


def max_feat_scatter(gfeats, max_idxs, bgs, lens, pn1):
    gfeats, max_idxs, bgs, lens = _f(gfeats), _i(max_idxs), _i(bgs), _i(lens)
    pn2, c = gfeats.shape
    out = np.zeros((pn1, c), np.float32)
    _lib.orc_max_feat_scatter(_p(gfeats), _p(max_idxs), _p(bgs), _p(lens), _i64(pn1), _i64(pn2), _i64(c), _p(out))
    return out


def locw_sum_backward(dlw_sum, lens, bgs, csum):
    dlw_sum, lens, bgs = _f(dlw_sum), _i(lens), _i(bgs)
    pn, m = dlw_sum.shape
    out = np.zeros((csum, m), np.float32)
    _lib.orc_locw_sum_backward(_p(dlw_sum), _p(lens), _p(bgs), _i64(pn), _i64(m), _i64(csum), _p(out))
    return out


def locw_feat_sum(lw, tfeats, lens, bgs):
    lw, tfeats, lens, bgs = _f(lw), _f(tfeats), _i(lens), _i(bgs)
    csum, m, ofn = tfeats.shape
    pn = lens.shape[0]
    out = np.zeros((pn, m, ofn), np.float64)
    _lib.orc_locw_feat_sum(_p(lw), _p(tfeats), _p(lens), _p(bgs), _i64(pn), _i64(m), _i64(ofn), _p(out))
    return out


def locw_feat_sum_backward(lw, tfeats, dsum, lens, bgs):
    """-> (dlw float64 [csum,m], dtfeats float32 [csum,m,ofn] (single multiply: bit-exact))."""
    lw, tfeats, dsum, lens, bgs = _f(lw), _f(tfeats), _f(dsum), _i(lens), _i(bgs)
    csum, m, ofn = tfeats.shape
    pn = lens.shape[0]
    dlw = np.zeros((csum, m), np.float64)
    dt = np.zeros((csum, m, ofn), np.float32)
    _lib.orc_locw_feat_sum_backward(_p(lw), _p(tfeats), _p(dsum), _p(lens), _p(bgs), _i64(pn), _i64(m), _i64(ofn),
                                    _i64(csum), _p(dlw), _p(dt))
    return dlw, dt


def eliminate_center(nidxs, lens, bgs):
    nidxs, lens, bgs = _i(nidxs), _i(lens), _i(bgs)
    pn, csum = lens.shape[0], nidxs.shape[0]
    on = np.zeros((csum - pn,), np.int32); oc = np.zeros((csum - pn,), np.int32)
    ol = np.zeros((pn,), np.int32); ob = np.zeros((pn,), np.int32)
    _lib.orc_eliminate_center(_p(nidxs), _p(lens), _p(bgs), _i64(pn), _p(on), _p(ol), _p(ob), _p(oc))
    return on, ol, ob, oc


# ---- pooling index ops -----------------------------------------------------------------------------
def voxel_index(pts, voxel_len, block_size):
    """a12; lower corner formed in fp32 as PoolingIndex/ComputeVoxelIdx.cc:66-67."""
    pts = _f(pts)
    f = np.float32
    mn_xy = -f(block_size) / f(2.0) - f(1e-3)
    mn_z = f(0.0) - f(1e-3)
    out = np.zeros((pts.shape[0], 3), np.int32)
    _lib.orc_voxel_index(_p(pts), _i64(pts.shape[0]), _i64(pts.shape[1]), C.c_float(float(f(voxel_len))),
                         C.c_float(float(mn_xy)), C.c_float(float(mn_xy)), C.c_float(float(mn_z)), _p(out))
    return out


def permutation_info(vidx, classes=None, block_ids=None):
    vidx = _i(vidx)
    pn = vidx.shape[0]
    classes = None if classes is None else _i(classes)
    block_ids = None if block_ids is None else _i(block_ids)
    o2p = np.zeros(pn, np.int32); p2o = np.zeros(pn, np.int32)
    lens = np.zeros(pn, np.int32); begs = np.zeros(pn, np.int32); cens = np.zeros(pn, np.int32)
    vn = _lib.orc_permutation_info(_p(vidx), _p(classes), _p(block_ids), _i64(pn), _p(o2p), _p(p2o), _p(lens),
                                   _p(begs), _p(cens))
    return o2p, p2o, lens[:vn].copy(), begs[:vn].copy(), cens


def diff_xyz(xyzs, cxyzs, cidxs):
    xyzs, cxyzs, cidxs = _f(xyzs), _f(cxyzs), _i(cidxs)
    out = np.zeros_like(xyzs)
    _lib.orc_diff_xyz(_p(xyzs), _p(cxyzs), _p(cidxs), _i64(xyzs.shape[0]), _p(out))
    return out


def repermutation_info(o2p2, lens, begs, pn1):
    o2p2, lens, begs = _i(o2p2), _i(lens), _i(begs)
    pn2 = o2p2.shape[0]
    ro2p = np.zeros(pn1, np.int32); rp2o = np.zeros(pn1, np.int32); rcens = np.zeros(pn1, np.int32)
    rlens = np.zeros(pn2, np.int32); rbegs = np.zeros(pn2, np.int32)
    _lib.orc_repermutation_info(_p(o2p2), _p(lens), _p(begs), _i64(pn1), _i64(pn2), _p(ro2p), _p(rp2o), _p(rlens),
                                _p(rbegs), _p(rcens))
    return ro2p, rp2o, rlens, rbegs, rcens


def _sq(x):
    return C.c_float(float(np.float32(x)))


def search(xyzs, sq_max, sq_min=None, block_begs=None):
    """a18/a19 -> (idxs, lens, begs, cens); sq_* are the squared radii as the TF attr carries them."""
    xyzs = _f(xyzs)
    pn = xyzs.shape[0]
    bb = None if block_begs is None else _i(block_begs)
    nb = 0 if bb is None else bb.shape[0] - 1
    use_min = C.c_int(0 if sq_min is None else 1)
    smin = _sq(0.0 if sq_min is None else sq_min)
    lens = np.zeros(pn, np.int32); begs = np.zeros(pn, np.int32)
    en = _lib.orc_search_count(_p(xyzs), _i64(pn), _p(bb), _i64(nb), smin, _sq(sq_max), use_min, _p(lens), _p(begs))
    idxs = np.zeros(en, np.int32); cens = np.zeros(en, np.int32)
    _lib.orc_search_fill(_p(xyzs), _i64(pn), _p(bb), _i64(nb), smin, _sq(sq_max), use_min, _p(begs), _p(idxs), _p(cens))
    return idxs, lens, begs, cens


def search_fixed(xyzs, sq_max, fixed_size, sq_min=None, block_begs=None):
    xyzs = _f(xyzs)
    pn, K = xyzs.shape[0], int(fixed_size)
    bb = None if block_begs is None else _i(block_begs)
    nb = 0 if bb is None else bb.shape[0] - 1
    use_min = C.c_int(0 if sq_min is None else 1)
    smin = _sq(0.0 if sq_min is None else sq_min)
    idxs = np.zeros(pn * K, np.int32); cens = np.zeros(pn * K, np.int32)
    lens = np.zeros(pn, np.int32); begs = np.zeros(pn, np.int32)
    _lib.orc_search_fixed(_p(xyzs), _i64(pn), _p(bb), _i64(nb), smin, _sq(sq_max), use_min, _i64(K), _p(idxs), _p(lens),
                          _p(begs), _p(cens))
    return idxs, lens, begs, cens


# ---- CSR helpers (tf_ops/test/data_util.py:12-27 compute_nidxs_bgs / compute_cidxs) -----------------
def bgs_from_lens(lens):
    lens = np.asarray(lens, np.int64)
    return (np.cumsum(lens) - lens).astype(np.int32)


def cidxs_from_lens(lens):
    return np.repeat(np.arange(len(lens), dtype=np.int32), np.asarray(lens, np.int64))

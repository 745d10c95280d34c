"""TEST INFRASTRUCTURE, NOT PRODUCT.  torch-facing ctypes wrapper of oracle/_ref/libref.so: the
reference's OWN unmodified CUDA sources (tf_ops/*.cu, tf_ops/PoolingIndex/*.cu) compiled for sm_100a
behind the extern "C" shim oracle/ref_shim.cu (built by oracle/Makefile in the authoring container;
the .so travels to the GPU box, /root/reference does not).  Used on the GPU box to pin the CPU oracle
and the product against the real reference, and by bench_ops.py as "the reference kernel on the same
B200" column.  The reference launches on the legacy default stream and cudaMalloc/cudaFree's inside
some calls; every wrapper synchronises afterwards.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

SO = Path(__file__).resolve().parent / "_ref" / "libref.so"
_lib = None


def available() -> bool:
    return SO.exists() and torch.cuda.is_available()


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(str(SO))
        _lib.ref_permutation_info.restype = C.c_int
        _lib.ref_search_count.restype = C.c_int
    return _lib


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _u(v):
    return C.c_uint(int(v))


def _f32(shape, dev):
    return torch.empty(shape, dtype=torch.float32, device=dev)


def _i32(shape, dev):
    return torch.empty(shape, dtype=torch.int32, device=dev)


SYNC = True  # False: rely on stream order (reference kernels and torch's default stream are both the legacy stream 0)


def _sync():
    if SYNC:
        lib().ref_sync()


def _pre():
    if SYNC:
        torch.cuda.synchronize()


def neighbor_scatter(ifeats, nidxs, lens, bgs, use_diff):
    _pre()
    pn, c = ifeats.shape
    out = _f32((nidxs.shape[0], c), ifeats.device)
    lib().ref_neighbor_scatter(_p(ifeats), _p(nidxs), _p(lens), _p(bgs), _u(pn), _u(c), _p(out), C.c_int(int(use_diff)))
    _sync()
    return out


def neighbor_gather(sfeats, nidxs, lens, bgs, use_diff):
    _pre()
    pn, c = lens.shape[0], sfeats.shape[1]
    out = _f32((pn, c), sfeats.device)
    lib().ref_neighbor_gather(_p(sfeats), _p(nidxs), _p(lens), _p(bgs), _u(pn), _u(c), _p(out), C.c_int(int(use_diff)))
    _sync()
    return out


def sum_feat_gather(feats, lens, bgs):
    _pre()
    pn, c = lens.shape[0], feats.shape[1]
    out = _f32((pn, c), feats.device)
    lib().ref_sum_feat_gather(_p(feats), _p(lens), _p(bgs), _u(pn), _u(c), _u(feats.shape[0]), _p(out))
    _sync()
    return out


def sum_feat_scatter(feats, cidxs):
    _pre()
    pn, c = feats.shape
    out = _f32((cidxs.shape[0], c), feats.device)
    lib().ref_sum_feat_scatter(_p(feats), _p(cidxs), _u(pn), _u(c), _u(cidxs.shape[0]), _p(out))
    _sync()
    return out


def max_feat_gather(feats, vlens, vbegs):
    _pre()
    pn2, c = vlens.shape[0], feats.shape[1]
    g = _f32((pn2, c), feats.device)
    mi = torch.zeros((pn2, c), dtype=torch.int32, device=feats.device)  # upstream leaves empty rows unset (N4)
    lib().ref_max_feat_gather(_p(feats), _p(vlens), _p(vbegs), _u(pn2), _u(c), _p(g), _p(mi))
    _sync()
    return g, mi


def max_feat_scatter(g, max_idxs, vbegs, vlens, pn1):
    _pre()
    pn2, c = g.shape
    out = _f32((pn1, c), g.device)
    lib().ref_max_feat_scatter(_p(g), _p(max_idxs), _p(vbegs), _p(vlens), _u(pn1), _u(pn2), _u(c), _p(out))
    _sync()
    return out


def locw_sum(lw, lens, bgs):
    _pre()
    pn, m = lens.shape[0], lw.shape[1]
    out = _f32((pn, m), lw.device)
    lib().ref_locw_sum(_p(lw), _p(lens), _p(bgs), _u(pn), _u(m), _p(out))
    _sync()
    return out


def locw_sum_backward(dsum, lens, bgs, csum):
    _pre()
    pn, m = dsum.shape
    out = _f32((csum, m), dsum.device)
    lib().ref_locw_sum_backward(_p(dsum), _p(lens), _p(bgs), _u(csum), _u(pn), _u(m), _p(out))
    _sync()
    return out


def locw_feat_sum(lw, tfeats, lens, bgs):
    _pre()
    csum, m, ofn = tfeats.shape
    pn = lens.shape[0]
    out = _f32((pn, m, ofn), lw.device)
    lib().ref_locw_feat_sum(_p(tfeats), _p(lw), _p(lens), _p(bgs), _u(pn), _u(ofn), _u(m), _p(out))
    _sync()
    return out


def locw_feat_sum_backward(lw, tfeats, dsum, lens, bgs):
    _pre()
    csum, m, ofn = tfeats.shape
    pn = lens.shape[0]
    dt = torch.zeros((csum, m, ofn), dtype=torch.float32, device=lw.device)
    dlw = _f32((csum, m), lw.device)
    lib().ref_locw_feat_sum_backward(_p(tfeats), _p(lw), _p(dsum), _p(lens), _p(bgs), _u(csum), _u(pn), _u(ofn), _u(m),
                                     _p(dt), _p(dlw))
    _sync()
    return dlw, dt


def eliminate_center(nidxs, lens, bgs):
    _pre()
    pn, csum = lens.shape[0], nidxs.shape[0]
    dev = nidxs.device
    on, oc, ol, ob = _i32((csum - pn,), dev), _i32((csum - pn,), dev), _i32((pn,), dev), _i32((pn,), dev)
    lib().ref_eliminate_center(_p(nidxs), _p(lens), _p(bgs), _u(pn), _p(on), _p(ol), _p(ob), _p(oc))
    _sync()
    return on, ol, ob, oc


def voxel_index(pts, voxel_len, mn_x, mn_y, mn_z):
    _pre()
    pn = pts.shape[0]
    out = _i32((pn, 3), pts.device)
    lib().ref_voxel_index(_p(pts), _p(out), C.c_int(pn), C.c_int(pts.shape[1]), C.c_float(voxel_len), C.c_float(mn_x),
                          C.c_float(mn_y), C.c_float(mn_z))
    _sync()
    return out


def permutation_info(vidx, classes=None):
    _pre()
    pn = vidx.shape[0]
    dev = vidx.device
    o2p, p2o, lens, begs, cens = (_i32((pn,), dev) for _ in range(5))
    vn = lib().ref_permutation_info(_p(vidx), _p(classes), _p(o2p), _p(p2o), _p(lens), _p(begs), _p(cens), C.c_int(pn))
    _sync()
    return o2p, p2o, lens[:vn].clone(), begs[:vn].clone(), cens


def repermutation_info(o2p2, lens, begs, pn1):
    _pre()
    pn2 = o2p2.shape[0]
    dev = o2p2.device
    rlens, rbegs = _i32((pn2,), dev), _i32((pn2,), dev)
    rcens, ro2p, rp2o = _i32((pn1,), dev), _i32((pn1,), dev), _i32((pn1,), dev)
    lib().ref_repermutation_info(_p(o2p2), _p(lens), _p(begs), _p(rlens), _p(rbegs), _p(rcens), _p(ro2p), _p(rp2o),
                                 C.c_int(pn1), C.c_int(pn2))
    _sync()
    return ro2p, rp2o, rlens, rbegs, rcens


def permutate_feature(feats, idxs):
    _pre()
    out = torch.empty_like(feats)
    lib().ref_permutate_feature(_p(feats), _p(idxs), _p(out), C.c_int(feats.shape[0]), C.c_int(feats.shape[1]))
    _sync()
    return out


def diff_xyz(xyzs, cxyzs, cidxs):
    _pre()
    out = torch.empty_like(xyzs)
    lib().ref_diff_xyz(_p(xyzs), _p(cxyzs), _p(out), _p(cidxs), C.c_int(xyzs.shape[0]), C.c_int(cxyzs.shape[0]))
    _sync()
    return out


def search(xyzs, sq_max, sq_min=None):
    _pre()
    pn = xyzs.shape[0]
    dev = xyzs.device
    use_min = C.c_int(0 if sq_min is None else 1)
    smin = C.c_float(0.0 if sq_min is None else sq_min)
    lens, begs = _i32((pn,), dev), _i32((pn,), dev)
    en = lib().ref_search_count(_p(xyzs), _p(lens), _p(begs), smin, C.c_float(sq_max), use_min, C.c_int(pn))
    idxs, cens = _i32((en,), dev), _i32((en,), dev)
    lib().ref_search_fill(_p(xyzs), _p(idxs), _p(cens), _p(begs), smin, C.c_float(sq_max), use_min, C.c_int(pn))
    _sync()
    return idxs, lens, begs, cens


def search_fixed(xyzs, sq_max, K, sq_min=None):
    _pre()
    pn = xyzs.shape[0]
    dev = xyzs.device
    use_min = C.c_int(0 if sq_min is None else 1)
    smin = C.c_float(0.0 if sq_min is None else sq_min)
    idxs, cens, lens, begs = _i32((pn * K,), dev), _i32((pn * K,), dev), _i32((pn,), dev), _i32((pn,), dev)
    lib().ref_search_fixed(_p(xyzs), _p(idxs), _p(cens), _p(lens), _p(begs), smin, C.c_float(sq_max), use_min,
                           C.c_int(K), C.c_int(pn))
    _sync()
    return idxs, lens, begs, cens

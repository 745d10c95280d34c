"""ctypes loader for libpcseg_ops.so — the C-ABI boundary declared in include/pcseg_ops.h.

Stands where the reference does `tf.load_op_library(path+'/build/libTFNeighborOps.so')`
(tf_ops/graph_conv_layer.py:4) and `.../PoolingOps.so` (tf_ops/graph_pooling_layer.py:4).
There is NO fallback: if the library is missing and cannot be built, import fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libpcseg_ops.so"

_p = C.c_void_p
_i64 = C.c_int64
_i = C.c_int
_f = C.c_float
_sz = C.c_size_t

# name -> (restype, argtypes); must list EVERY symbol of include/pcseg_ops.h (tests check this).
SIGNATURES: dict[str, tuple] = {
    "pcs_last_error": (C.c_char_p, []),
    "pcs_version": (_i, []),
    "pcs_launch_count": (_i64, []),
    "pcs_neighbor_scatter": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p]),
    "pcs_csr_cidxs": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "pcs_csr_transpose_workspace_bytes": (_sz, [_i64, _i64]),
    "pcs_csr_transpose": (_i, [_p, _i64, _i64, _p, _p, _p, _sz, _p]),
    "pcs_csr_transpose_symmetric": (_i, [_p, _p, _p, _p, _i64, _i64, _p, _p, _p, _p]),
    "pcs_neighbor_gather": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p]),
    "pcs_neighbor_sum_feat_gather": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_neighbor_sum_feat_scatter": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_neighbor_max_feat_gather": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p]),
    "pcs_neighbor_max_feat_scatter": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_sum": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_sum_backward": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_feat_sum": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_feat_sum_backward": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p]),
    "pcs_eliminate_center": (_i, [_p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p]),
    "pcs_compute_voxel_index": (_i, [_p, _i64, _i64, _f, _f, _f, _f, _p, _p]),
    "pcs_compute_permutation_info_workspace_bytes": (_sz, [_i64]),
    "pcs_compute_permutation_info": (_i, [_p, _p, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_compute_permutation_info64": (_i, [_p, _p, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_permutate_feature": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "pcs_compute_diff_xyz": (_i, [_p, _p, _p, _i64, _i64, _p, _p]),
    "pcs_compute_repermutation_info_workspace_bytes": (_sz, [_i64]),
    "pcs_compute_repermutation_info": (_i, [_p, _p, _p, _i64, _i64, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_search_workspace_bytes": (_sz, [_i64]),
    "pcs_search_neighborhood_count": (_i, [_p, _i64, _p, _i64, _i64, _f, _f, _i, _p, _p, _p, _p, _sz, _p]),
    "pcs_search_neighborhood_fill": (_i, [_p, _i64, _p, _i64, _i64, _f, _f, _i, _p, _p, _p, _p, _sz, _p]),
    "pcs_search_neighborhood_fixed": (_i, [_p, _i64, _p, _i64, _i64, _f, _f, _i, _i64, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_dense_mlp3_supported": (_i, [_i64, _i64, _i64, _i64]),
    "pcs_dense_mlp3_forward": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p]),
    "pcs_dense_mlp3_backward": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _i64, _p, _p, _p]),
    "pcs_set_option": (_i, [C.c_char_p, _i64]),
    "pcs_dense_mlp3_backward_pooled": (_i, [_p] * 8 + [_i64] * 5 + [_p, _p, _p, _p]),
    "pcs_dense_mlp3_wgrad_partials": (_i64, [_i64] * 5),
    "pcs_pool_linear_wgrad_partials": (_i64, [_i64, _i64]),
    "pcs_pool_linear_wgrad": (_i, [_p] * 5 + [_i64] * 4 + [_p, _p]),
    "pcs_column_sum_partials": (_i64, [_i64, _i64]),
    "pcs_column_sum": (_i, [_p, _i64, _i64, _p, _p, _p]),
    "pcs_edge_combine_forward": (_i, [_p] * 8 + [_i64] * 4 + [_p, _p]),
    "pcs_edge_xyz_wgrad_partials": (_i64, [_i64, _i64]),
    "pcs_edge_xyz_wgrad": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "pcs_linear_tc_workspace_bytes": (_sz, [_i64, _i64]),
    "pcs_linear_tc": (_i, [_p, _i64, _i64, _i64, _p, _i64, _i64, _i64, _p, _p, _i64, _i, _p, _i64, _p, _sz, _p]),
    "pcs_linear_tc_wgrad_partials": (_i64, [_i64, _i64, _i64, _i]),
    "pcs_linear_tc_wgrad": (_i, [_p, _i64, _i64, _i64, _p, _i64, _i64, _i, _p, _p]),
    "pcs_relu_concat_dropout_forward": (_i, [_p, _i64, _i64, _i64, _p, _i64, _i64, _f, C.c_uint64, _p, _p]),
    "pcs_relu_concat_dropout_backward": (_i, [_p, _p, _i64, _i64, _i64, _f, C.c_uint64, _p, _i64, _p, _i64, _p]),
    "pcs_neighbor_scatter_bf16": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p]),
    "pcs_neighbor_gather_bf16": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i, _p, _p]),
    "pcs_neighbor_sum_feat_gather_bf16": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_neighbor_sum_feat_scatter_bf16": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_neighbor_max_feat_gather_bf16": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p, _p]),
    "pcs_neighbor_max_feat_scatter_bf16": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_sum_bf16": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_sum_backward_bf16": (_i, [_p, _p, _p, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_feat_sum_bf16": (_i, [_p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p]),
    "pcs_location_weight_feat_sum_backward_bf16": (_i, [_p, _p, _p, _p, _p, _i64, _i64, _i64, _i64, _p, _p, _p]),
    "pcs_permutate_feature_bf16": (_i, [_p, _p, _i64, _i64, _p, _p]),
    "pcs_search_multi_count": (_i, [_p, _i64, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_search_multi_fill": (_i, [_p, _i64, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "pcs_search_multi_count_pad": (_i, [_p, _i64, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p, _i64, _p, _sz, _p]),
    "pcs_search_multi_fill_pad": (_i, [_p, _i64, _p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _i64, _p, _p, _p, _sz, _p]),
    "pcs_cross_search_workspace_bytes": (_sz, [_i64, _i64]),
    "pcs_cross_search_count": (_i, [_p, _i64, _p, _i64, _f, _f, _i, _p, _p, _p, _p, _sz, _p]),
    "pcs_cross_search_fill": (_i, [_p, _i64, _p, _i64, _f, _f, _i, _p, _p, _p, _p, _sz, _p]),
    "pcs_knn_cross": (_i, [_p, _i64, _p, _i64, _i64, _f, _i64, _p, _p, _p, _p, _sz, _p]),
    "pcs_interpolate_probs": (_i, [_p, _p, _p, _p, _p, _p, _i64, _i64, _f, _p, _p]),
    "pcs_compute_covars": (_i, [_p, _p, _p, _p, _i64, _p, _p]),
    "pcs_block_members_tiles": (_i64, [_i64]),
    "pcs_block_members_count": (_i, [_p, _i64, _p, _i64, _f, _p, _p]),
    "pcs_block_members_fill": (_i, [_p, _i64, _p, _i64, _f, _p, _p, _p]),
    "pcs_exclusive_scan_workspace_bytes": (_sz, [_i64]),
    "pcs_exclusive_scan_i32": (_i, [_p, _i64, _p, _p, _p, _sz, _p]),
    "pcs_exclusive_scan_static_max_n": (_i64, []),
    "pcs_exclusive_scan_i32_static": (_i, [_p, _i64, _p, _p, _p]),
}


# host-only queries: no status code, no launch, never profiled
_PLAIN = ("pcs_last_error", "pcs_version", "pcs_launch_count", "pcs_dense_mlp3_supported", "pcs_edge_xyz_wgrad_partials",
          "pcs_pool_linear_wgrad_partials", "pcs_set_option", "pcs_exclusive_scan_static_max_n", "pcs_block_members_tiles",
          "pcs_dense_mlp3_wgrad_partials", "pcs_column_sum_partials", "pcs_linear_tc_wgrad_partials")


class PcsError(RuntimeError):
    """A libpcseg_ops entry point returned non-zero (the reference would print and exit(),
    tf_ops/TFCudaCommon.h:14-22)."""


def _load() -> C.CDLL:
    if os.environ.get("PCSEG_NO_AUTOBUILD"):
        if not LIB_PATH.exists():
            raise ImportError(f"{LIB_PATH} is missing (PCSEG_NO_AUTOBUILD set); there is no CPU fallback")
    else:
        from . import build as _build  # nvcc cross-compiles without a GPU

        if not LIB_PATH.exists() or Path(_build.NVCC).exists():
            _build.build()  # content-hashed: a no-op unless csrc/ or the header changed since the last build
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    return lib


_clib = _load()


class _Profile:
    """Optional per-call device timing: when enabled, every C-ABI compute call is bracketed by CUDA
    events on the caller's current stream (the stream the kernels are launched on).  Off by default;
    bench.py turns it on to attribute step time to kernels and compute achieved GB/s."""

    enabled = False
    records: list = []  # (entry name, integer/float args, start event, end event)


profile = _Profile()


class _Lib:
    """Attribute proxy over the CDLL so that calls can be timed without touching the op code."""

    def __getattr__(self, name):
        fn = getattr(_clib, name)
        if name.endswith("_bytes") or name in _PLAIN:
            object.__setattr__(self, name, fn)
            return fn

        def call(*args):
            if not profile.enabled:
                return fn(*args)
            import torch

            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            st = fn(*args)
            e1.record()
            profile.records.append((name, args, e0, e1))
            return st

        object.__setattr__(self, name, call)
        return call


lib = _Lib()


def last_error() -> str:
    return lib.pcs_last_error().decode(errors="replace")


def launch_count() -> int:
    return int(lib.pcs_launch_count())


def check(status: int, what: str) -> None:
    if status != 0:
        raise PcsError(f"{what}: {last_error()}")

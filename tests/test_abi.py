"""CPU: the C-ABI library loads and exports every symbol include/pcseg_ops.h declares (no compute
calls — there is no GPU here), the ctypes table covers the header, and the host layer validates
arguments like the reference's OP_REQUIRES."""
from __future__ import annotations

import ctypes
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
HEADER = (ROOT / "include" / "pcseg_ops.h").read_text()


def header_symbols():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(pcs_[a-z0-9_]+)\s*\(", body)))


def test_header_has_symbols():
    syms = header_symbols()
    assert len(syms) >= 25 and "pcs_neighbor_scatter" in syms


def test_library_exports_every_header_symbol():
    from pointcloudsegmentation_b200 import LIB_PATH

    lib = ctypes.CDLL(str(LIB_PATH))
    for s in header_symbols():
        assert hasattr(lib, s), f"{s} declared in pcseg_ops.h but not exported"


def test_ctypes_table_matches_header():
    from pointcloudsegmentation_b200._lib import SIGNATURES

    assert sorted(SIGNATURES) == header_symbols()


def test_sizes_and_version_callable_without_gpu():
    from pointcloudsegmentation_b200._lib import lib

    assert lib.pcs_version() >= 100
    assert lib.pcs_launch_count() >= 0
    assert lib.pcs_compute_permutation_info_workspace_bytes(1000) > 1000 * 8
    assert lib.pcs_csr_transpose_workspace_bytes(10, 100) >= 800
    assert lib.pcs_search_workspace_bytes(100) > 0


def test_header_cites_reference_for_every_op():
    # every op block names the reference launcher it replaces
    assert HEADER.count("replaces:") >= 18


def test_no_cpu_path():
    """CPU tensors are rejected loudly — there is no fallback (reference ops are DEVICE_GPU only)."""
    from pointcloudsegmentation_b200 import neighbor_ops

    f = torch.zeros(4, 3)
    i = torch.zeros(4, dtype=torch.int32)
    with pytest.raises(ValueError, match="CUDA"):
        neighbor_ops.neighbor_scatter(f, i, i, i, use_diff=False)


def test_product_never_imports_oracle():
    pkg = ROOT / "pointcloudsegmentation_b200"
    for p in pkg.rglob("*.py"):
        src = p.read_text()
        assert "oracle" not in src.replace("# oracle", ""), f"{p} mentions the oracle"
    for p in list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")):
        assert "oracle/" not in p.read_text()

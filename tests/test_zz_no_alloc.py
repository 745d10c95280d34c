"""pcs_exclusive_scan_i32_static — the allocation-free twin of pcs_exclusive_scan_i32 (scratch in a file-scope
`static __device__` array, csrc/scan_static.cu).  No counterpart in the reference.
Collected last on purpose.

Checked through the C ABI: bit-exact against the existing entry point (int32 work) on two calls in a row with
different inputs (the shared array keeps the previous call's contents), at the edge sizes, and rejected beyond
the stated maximum with nothing written.
"""
from __future__ import annotations

import re
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def _both(x):
    """(static out, static total, workspace out, workspace total) for one int32 device vector."""
    from pointcloudsegmentation_b200._lib import check, lib
    from pointcloudsegmentation_b200.ops import _ptr, _stream

    n = x.numel()
    out_s = torch.full((n,), -7, dtype=torch.int32, device=x.device)
    tot_s = torch.full((1,), -7, dtype=torch.int32, device=x.device)
    check(lib.pcs_exclusive_scan_i32_static(_ptr(x), n, _ptr(out_s), _ptr(tot_s), _stream()), "scan_static")
    out_w = torch.full((n,), -9, dtype=torch.int32, device=x.device)
    tot_w = torch.full((1,), -9, dtype=torch.int32, device=x.device)
    ws = torch.empty(lib.pcs_exclusive_scan_workspace_bytes(n), dtype=torch.uint8, device=x.device)
    check(lib.pcs_exclusive_scan_i32(_ptr(x), n, _ptr(out_w), _ptr(tot_w), _ptr(ws), ws.numel(), _stream()), "scan")
    torch.cuda.synchronize()
    return out_s, tot_s, out_w, tot_w


@pytest.mark.gpu
def test_static_scan_matches_workspace_scan_twice_in_a_row():
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(20261105)
    # two calls in a row on the same shared scratch: a large one first, then a smaller one with other values
    for n, hi in ((1_000_003, 40), (515_211, 27)):
        x_np = rng.integers(0, hi, n).astype(np.int32)
        x = torch.from_numpy(x_np).to(dev)
        out_s, tot_s, out_w, tot_w = _both(x)
        assert torch.equal(out_s, out_w)  # integer work: bit-exact
        assert torch.equal(tot_s, tot_w)
        want = np.cumsum(x_np, dtype=np.int64) - x_np
        assert np.array_equal(out_s.cpu().numpy().astype(np.int64), want)
        assert int(tot_s.item()) == int(x_np.sum(dtype=np.int64))


@pytest.mark.gpu
def test_static_scan_edge_sizes_and_limit():
    from pointcloudsegmentation_b200._lib import lib, last_error
    from pointcloudsegmentation_b200.ops import _ptr, _stream

    dev = torch.device("cuda:0")
    max_n = lib.pcs_exclusive_scan_static_max_n()
    assert max_n == 1 << 22
    for n in (0, 1, 31, 4096, max_n):
        x = (torch.arange(n, dtype=torch.int32, device=dev) % 5)
        out_s, tot_s, out_w, tot_w = _both(x)
        assert torch.equal(out_s, out_w) and torch.equal(tot_s, tot_w)
    # total_dev is nullable
    x = torch.ones(100, dtype=torch.int32, device=dev)
    out = torch.empty_like(x)
    assert lib.pcs_exclusive_scan_i32_static(_ptr(x), 100, _ptr(out), None, _stream()) == 0
    torch.cuda.synchronize()
    assert torch.equal(out, torch.arange(100, dtype=torch.int32, device=dev))
    # beyond the maximum: refused with a clear message, nothing launched, the output untouched
    n = max_n + 1
    x = torch.ones(n, dtype=torch.int32, device=dev)
    out = torch.full((n,), -3, dtype=torch.int32, device=dev)
    before = lib.pcs_launch_count()
    assert lib.pcs_exclusive_scan_i32_static(_ptr(x), n, _ptr(out), None, _stream()) != 0
    assert "exceeds the static scratch limit" in last_error()
    assert lib.pcs_launch_count() == before
    torch.cuda.synchronize()
    assert bool((out == -3).all())


def test_static_scan_source_allocates_nothing():
    """CPU check of the contract: the translation unit names no allocating routine and takes the array's device
    address through cudaGetSymbolAddress (the symbol itself is only a host-side stand-in)."""
    src = (ROOT / "pointcloudsegmentation_b200" / "csrc" / "scan_static.cu").read_text()
    code = re.sub(r"//.*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaMallocAsync", "cudaHostAlloc", "thrust::", "new "):
        assert banned not in code, banned
    assert "static __device__ __align__(256)" in code
    assert "cudaGetSymbolAddress" in code

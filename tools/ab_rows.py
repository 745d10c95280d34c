#!/usr/bin/env python
"""A/B of alternative kernels for the same row-routing op (development aid; feeds the dispatch
heuristics in csrc/neighbor_ops.cu / neighbor_pool.cu).  Same workload and timer as bench_ops.py."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from bench_ops import Timer  # noqa: E402
from pointcloudsegmentation_b200 import ops as P  # noqa: E402
from pointcloudsegmentation_b200._lib import check, lib  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_crop  # noqa: E402

B, pn = 64, 4096
rng = np.random.default_rng(0)
pts = torch.from_numpy(np.concatenate([s3dis_crop(rng, pn) for _ in range(B)])).cuda()
bb = torch.from_numpy((np.arange(B + 1) * pn).astype(np.int32)).cuda()
nidxs, nlens, nbegs, ncens = P.search_neighborhood_brute_force(pts, 0.01, block_begs=bb, max_block_pn=pn)
Pn, E = B * pn, nidxs.shape[0]
T = Timer(10)
st = torch.cuda.current_stream().cuda_stream
peak = 6555.2


def scatter(f, cidxs, diff, out):
    check(lib.pcs_neighbor_scatter(f.data_ptr(), nidxs.data_ptr(), nlens.data_ptr(), nbegs.data_ptr(),
                                   cidxs.data_ptr() if cidxs is not None else None, Pn, f.shape[1], E, diff,
                                   out.data_ptr(), st), "scatter")


for C in [int(c) for c in (sys.argv[1] if len(sys.argv) > 1 else "3,12,32,35,64,67,128,131").split(",")]:
    f = torch.rand(Pn, C, device="cuda")
    out = torch.empty(E, C, device="cuda")
    by = 4 * C * (Pn + E) + 4 * E + 8 * Pn
    r = {}
    for o in ("diff", "diff_odd", "copy", "copy_odd"):
        lib.pcs_set_option(f"scatter_walk_min_c_{o}".encode(), -1)
    r["scatter d0 flat"] = T(lambda: scatter(f, None, 0, out))
    r["scatter d1 flat(cidxs)"] = T(lambda: scatter(f, ncens, 1, out))
    for o in ("diff", "diff_odd", "copy", "copy_odd"):
        lib.pcs_set_option(f"scatter_walk_min_c_{o}".encode(), 0)
    r["scatter d0 walk"] = T(lambda: scatter(f, None, 0, out))
    r["scatter d1 walk"] = T(lambda: scatter(f, None, 1, out))
    r["sum_feat_scatter flat"] = T(lambda: P.neighbor_sum_feat_scatter(f, ncens, nlens, nbegs))
    r["row_broadcast (a8 kernel)"] = T(lambda: P.location_weight_sum_backward(out, f, nlens, nbegs))
    print(f"C={C}: " + "  ".join(f"{k} {1e3 * v:.1f}us ({100 * by / 1e9 / (v * 1e-3) / peak:.0f}%)" for k, v in r.items()), flush=True)

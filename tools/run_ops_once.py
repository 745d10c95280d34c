#!/usr/bin/env python
"""Launch each selected op once (after one warm-up) on the op-level batch — the short program `ncu --set full` wraps.

    python tools/run_ops_once.py --ops scatter,scatter_diff,gather,gather_diff,sum_gather,max_gather,permutate --channels 3,35,131
"""
import argparse
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
ap = argparse.ArgumentParser()
ap.add_argument("--ops", default="scatter,scatter_diff,gather,gather_diff,sum_gather,max_gather,sum_scatter,max_scatter,permutate")
ap.add_argument("--channels", default="3,35,131")
ap.add_argument("--blocks", type=int, default=128)
ap.add_argument("--dtype", default="f32")
a = ap.parse_args()

from pointcloudsegmentation_b200 import ops as P  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_crop  # noqa: E402

rng = np.random.default_rng(0)
B, pn = a.blocks, 4096
pts = np.tile(np.concatenate([s3dis_crop(rng, pn) for _ in range(16)]), (B // 16, 1))
raw = torch.from_numpy(pts).cuda()
bids = torch.arange(B, device="cuda", dtype=torch.int32).repeat_interleave(pn)
bb = (torch.arange(B + 1, device="cuda") * pn).to(torch.int32)
o2p = P.compute_permutation_info(P.compute_voxel_index(raw, 0.15, 3.0), block_ids=bids)[0]
x = P.permutate_feature(raw, o2p)
ni, nl, nb, nc = P.search_neighborhood_brute_force(x, 0.01, block_begs=bb, max_block_pn=pn)
P.csr_transpose(ni, B * pn)
E = ni.shape[0]
dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
want = set(a.ops.split(","))
for C in [int(c) for c in a.channels.split(",")]:
    fp = (torch.rand(B * pn, C, device="cuda") * 2 - 1).to(dt)
    fe = (torch.rand(E, C, device="cuda") * 2 - 1).to(dt)
    g, mi = P.neighbor_max_feat_gather(fe, nl, nb)
    table = {
        "scatter": lambda: P.neighbor_scatter(fp, ni, nl, nb, False), "scatter_diff": lambda: P.neighbor_scatter(fp, ni, nl, nb, True),
        "gather": lambda: P.neighbor_gather(fe, ni, nl, nb, False), "gather_diff": lambda: P.neighbor_gather(fe, ni, nl, nb, True),
        "sum_gather": lambda: P.neighbor_sum_feat_gather(fe, nc, nl, nb), "max_gather": lambda: P.neighbor_max_feat_gather(fe, nl, nb),
        "sum_scatter": lambda: P.neighbor_sum_feat_scatter(fp, nc, nl, nb), "max_scatter": lambda: P.neighbor_max_feat_scatter(g, E, mi, nb, nl),
        "permutate": lambda: P.permutate_feature(fp, o2p), "eliminate": lambda: P.eliminate_center(ni, nl, nb),
    }
    for name, fn in table.items():
        if name in want:
            fn(); torch.cuda.synchronize()
            torch.cuda.nvtx.range_push(f"{name}_C{C}")
            fn(); torch.cuda.synchronize()
            torch.cuda.nvtx.range_pop()
    del fp, fe, g, mi
print("ok", E)

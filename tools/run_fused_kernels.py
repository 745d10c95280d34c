#!/usr/bin/env python
"""Launch every fused edge-conv kernel once at S3DIS level-0 / level-1 sizes (ncu target; no timing)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200 import fused_ops, ops as P  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_crop  # noqa: E402

B, pn = 32, 4096
rng = np.random.default_rng(0)
pts = torch.from_numpy(np.concatenate([s3dis_crop(rng, pn) for _ in range(B)])).cuda()
bb = torch.from_numpy((np.arange(B + 1) * pn).astype(np.int32)).cuda()
nidxs, nlens, nbegs, ncens = P.search_neighborhood_brute_force(pts, 0.01, block_begs=bb, max_block_pn=pn)
E = nidxs.shape[0]
sx = P.neighbor_scatter(pts, nidxs, nlens, nbegs, True)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for A, Bq, C, O in [(8, 8, 16, 32), (16, 16, 16, 48), (16, 16, 32, 64)]:
    H = A + Bq + C + O
    g = torch.Generator(device="cuda").manual_seed(0)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    pab = r(B * pn, 2 * H).requires_grad_(True)
    wx = r(H, 3).requires_grad_(True)
    w2, w3, wo = (t.requires_grad_(True) for t in (r(Bq, A) / 3, r(C, A + Bq) / 4, r(O, A + Bq + C) / 6))
    for _ in range(reps):
        u = fused_ops.edge_combine(pab, sx, wx, nidxs, nlens, nbegs, ncens)
        y = fused_ops.dense_mlp3_pool(u, w2, w3, wo, nlens, nbegs, ncens)[0]
        y.backward(torch.ones_like(y))
    torch.cuda.synchronize()
    print(f"cfg ({A},{Bq},{C},{O}): E={E} rows, H={H}")

#!/usr/bin/env python
"""Per-call integer arguments and device times of one C-ABI entry inside one 64-block S3DIS train step (CUDA events around
every call, _lib.profile).  usage: op_shapes.py pcs_pool_linear_wgrad [more names]"""
import sys
from collections import defaultdict
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200._lib import profile  # noqa: E402
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget  # noqa: E402

names = sys.argv[1:] or ["pcs_pool_linear_wgrad"]
B = 64
d = s3dis_batch_to_budget(0, B, B * 8072)
b = {k: torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids")}
torch.manual_seed(0)
m = Pointnet13DilatedEmbed().cuda().train()
with torch.no_grad():
    m(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
opt = torch.optim.Adam(m.parameters(), lr=1e-3, fused=True)
prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)


def step():
    opt.zero_grad(set_to_none=True)
    loss, _, _ = m(prepared=prep)
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
profile.records.clear()
profile.enabled = True
step()
torch.cuda.synchronize()
profile.enabled = False
for name in names:
    agg = defaultdict(lambda: [0, 0.0])
    for n, a, e0, e1 in profile.records:
        if n != name:
            continue
        key = tuple(int(v) for v in a if isinstance(v, int) and not isinstance(v, bool) and abs(int(v)) < (1 << 40))
        agg[key][0] += 1
        agg[key][1] += e0.elapsed_time(e1) * 1e3
    tot = sum(v[1] for v in agg.values())
    print(f"{name}: total {tot / 1e3:.2f} ms")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  x{v[0]:2d} {v[1] / v[0]:9.1f} us/call  args {k}")

#!/usr/bin/env python
"""Per-call shapes and device times of pcs_linear_tc / pcs_linear_tc_wgrad inside one 64-block S3DIS train step
(CUDA events around every C-ABI call, _lib.profile): which layers the GEMM time of a step is made of."""
import sys
from collections import defaultdict
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200._lib import profile  # noqa: E402
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget  # noqa: E402

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
agg = defaultdict(lambda: [0, 0.0])
for name, a, e0, e1 in profile.records:
    if name == "pcs_linear_tc":
        key = (name, int(a[1]), int(a[2]), int(a[5]), bool(a[9]))  # M, K, N, addend
    elif name == "pcs_linear_tc_wgrad":
        key = (name, int(a[1]), int(a[2]), int(a[5]), False)
    else:
        continue
    agg[key][0] += 1
    agg[key][1] += e0.elapsed_time(e1) * 1e3
tot = sum(v[1] for v in agg.values())
print(f"total {tot / 1e3:.2f} ms")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    name, M, K, N, add = k
    byt = 4 * (M * K + M * N * (2 if add else 1) + N * K)
    us = v[1] / v[0]
    print(f"{name[4:]:16s} M={M:7d} K={K:5d} N={N:5d} add={int(add)} x{v[0]:2d} {us:8.1f} us/call  {byt / us / 1e3:7.1f} GB/s  {2 * M * K * N / us / 1e6:7.1f} TFLOP/s  total {v[1] / 1e3:6.2f} ms")

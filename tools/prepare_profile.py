#!/usr/bin/env python
"""Device time of every C-ABI call inside one prepare() of the 64-block S3DIS batch, run ALONE on the GPU."""
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
m = Pointnet13DilatedEmbed().cuda()
for _ in range(3):
    m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
torch.cuda.synchronize()
profile.records.clear()
profile.enabled = True
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
e1.record()
torch.cuda.synchronize()
profile.enabled = False
agg = defaultdict(lambda: [0, 0.0])
for n, a, s0, s1 in profile.records:
    agg[n][0] += 1
    agg[n][1] += s0.elapsed_time(s1) * 1e3
print(f"prepare: {e0.elapsed_time(e1):.2f} ms on the device, C-ABI calls {sum(v[1] for v in agg.values()) / 1e3:.2f} ms")
for n, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {n:40s} x{v[0]:3d} {v[1]:9.1f} us")

#!/usr/bin/env python
"""Device time of every C-ABI entry inside one training step (prepare excluded) of the 64-block S3DIS batch, plus the time
between the entries (stock torch kernels)."""
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
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); step(); e1.record(); torch.cuda.synchronize()
plain = e0.elapsed_time(e1)
profile.records.clear()
profile.enabled = True
e0.record(); step(); e1.record()
torch.cuda.synchronize()
profile.enabled = False
agg = defaultdict(lambda: [0, 0.0])
for n, a, s0, s1 in profile.records:
    agg[n][0] += 1
    agg[n][1] += s0.elapsed_time(s1)
own = sum(v[1] for v in agg.values())
print(f"step {plain:.2f} ms (with events {e0.elapsed_time(e1):.2f}); C-ABI entries {own:.2f} ms in {sum(v[0] for v in agg.values())} calls")
for n, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {n:42s} x{v[0]:4d} {v[1]:8.3f} ms")

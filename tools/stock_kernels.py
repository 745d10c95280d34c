#!/usr/bin/env python
"""CUDA kernels of one training step grouped by name (torch.profiler / CUPTI): which stock torch kernels are left."""
import sys
from pathlib import Path

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
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
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
rows = [e for e in prof.key_averages() if e.device_time_total > 0 and e.device_type.name == "CUDA"]
rows.sort(key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in rows)
print(f"kernels: {tot / 1e3:.2f} ms")
for e in rows[:45]:
    print(f"{e.device_time_total / 1e3:8.3f} ms x{e.count:4d}  {e.key[:120]}")

print()
ops = [e for e in prof.key_averages(group_by_input_shape=True) if e.key.startswith("aten::") and e.device_time_total > 50]
ops.sort(key=lambda e: -e.device_time_total)
for e in ops[:40]:
    print(f"{e.device_time_total / 1e3:8.3f} ms x{e.count:4d}  {e.key:28s} {str(e.input_shapes)[:110]}")

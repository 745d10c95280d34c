#!/usr/bin/env python
"""Which reference cycles does one training step leave behind (objects only the cyclic collector frees), and do they hold
device tensors?"""
import gc
import sys
from collections import Counter
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget  # noqa: E402

B = 8
d = s3dis_batch_to_budget(0, B, B * 8072)
b = {k: torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids")}
m = Pointnet13DilatedEmbed().cuda().train()
with torch.no_grad():
    m(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
opt = torch.optim.Adam(m.parameters(), lr=1e-3, fused=True)


def step():
    opt.zero_grad(set_to_none=True)
    prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
    loss, _, _ = m(prepared=prep)
    loss.backward()
    opt.step()


for _ in range(2):
    step()
torch.cuda.synchronize()
gc.collect()
gc.disable()
m0 = torch.cuda.memory_allocated()
step()
torch.cuda.synchronize()
m1 = torch.cuda.memory_allocated()
gc.set_debug(gc.DEBUG_SAVEALL)
n = gc.collect()
m2 = torch.cuda.memory_allocated()
print(f"unreachable objects after one step: {n}; device memory {m0 >> 20} -> {m1 >> 20} MiB after the step, {m2 >> 20} MiB after collect")
c = Counter(type(o).__name__ for o in gc.garbage)
print(c.most_common(15))
tens = [o for o in gc.garbage if torch.is_tensor(o)]
print("tensors in cycles:", len(tens), sum(t.numel() * t.element_size() for t in tens if t.is_cuda) >> 20, "MiB")
for o in gc.garbage:
    if type(o).__name__ in ("function", "cell", "method") :
        print("  ", type(o).__name__, getattr(o, "__qualname__", ""), str(o)[:100])
        break
fn = [o for o in gc.garbage if type(o).__name__.endswith("Backward") or "Backward" in type(o).__name__]
print(Counter(type(o).__name__ for o in fn).most_common(10))

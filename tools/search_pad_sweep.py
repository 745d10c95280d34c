#!/usr/bin/env python
"""Degree distribution of the tower's radius graphs on the bench workload, and the prepare time per padded-list length."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200 import ops  # noqa: E402
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget  # noqa: E402

B = 64
d = s3dis_batch_to_budget(0, B, B * 8072)
b = {k: torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids")}
m = Pointnet13DilatedEmbed().cuda()
for K in (0, 32, 64, 96, 128, 192, 256):
    ops.SEARCH_PAD_K = K
    for _ in range(2):
        prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
    torch.cuda.synchronize()
    print(f"pad_k {K:4d}: prepare {(time.perf_counter() - t0) / 5 * 1e3:7.2f} ms")
graphs = prep[-2] if isinstance(prep, (list, tuple)) else None
for name in dir(prep) if graphs is None else []:
    pass
# degree statistics of every graph the tower builds
import itertools
def walk(x):
    if torch.is_tensor(x):
        yield x
    elif isinstance(x, (list, tuple)):
        for y in x:
            yield from walk(y)
    elif isinstance(x, dict):
        for y in x.values():
            yield from walk(y)
seen = set()
for t in walk(prep):
    if t.dtype == torch.int32 and t.dim() == 1 and t.numel() in (516607, 94916, 12332) and t.data_ptr() not in seen and int(t.max()) < 4096 and int(t.min()) >= 0 and float(t.float().mean()) > 1.5:
        seen.add(t.data_ptr())
        q = torch.quantile(t.float()[:: max(1, t.numel() // 200000)], torch.tensor([0.5, 0.9, 0.99], device="cuda"))
        print(f"lens-like [{t.numel()}]: mean {float(t.float().mean()):6.1f} p50 {q[0]:.0f} p90 {q[1]:.0f} p99 {q[2]:.0f} max {int(t.max())}  frac>64 {float((t > 64).float().mean()):.4f} frac>128 {float((t > 128).float().mean()):.4f}")

#!/usr/bin/env python
"""Host cost of model.prepare (voxel pooling + searches + transposes) in isolation."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch  # noqa: E402

B = 16
d = s3dis_batch(0, B)
b = {k: torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids")}
torch.manual_seed(0)
m = Pointnet13DilatedEmbed().cuda()
m.train()
for _ in range(3):
    prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"prepare wall ms (host, incl. its syncs, idle GPU): {(t1 - t0) * 100:.2f}")
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)

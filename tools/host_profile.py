#!/usr/bin/env python
"""Where does the HOST time of a training step go?  cProfile over a few steps of the S3DIS tower (main thread:
forward + optimizer; the autograd engine's backward thread is profiled through torch.profiler's CPU table)."""
import cProfile
import pstats
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed  # noqa: E402
from pointcloudsegmentation_b200.synthetic import s3dis_batch  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
d = s3dis_batch(0, B)
b = {k: torch.from_numpy(d[k]).cuda() for k in ("xyzs", "feats", "labels", "block_ids")}
torch.manual_seed(0)
m = Pointnet13DilatedEmbed().cuda()
with torch.no_grad():
    m(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
opt = torch.optim.Adam(m.parameters(), lr=1e-3, fused=True)
prep = m.prepare(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)


def fwd():
    opt.zero_grad(set_to_none=True)
    loss, _, _ = m(prepared=prep)
    return loss


def step():
    loss = fwd()
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
N = 5
t = [0.0, 0.0, 0.0]
for _ in range(N):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    loss = fwd()
    t1 = time.perf_counter()
    loss.backward()
    t2 = time.perf_counter()
    opt.step()
    t3 = time.perf_counter()
    t[0] += t1 - t0
    t[1] += t2 - t1
    t[2] += t3 - t2
    torch.cuda.synchronize()
print(f"host ms/step (GPU idle at start): forward {t[0] / N * 1e3:.2f}  backward {t[1] / N * 1e3:.2f}  adam {t[2] / N * 1e3:.2f}")
pr = cProfile.Profile()
pr.enable()
for _ in range(N):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(35)

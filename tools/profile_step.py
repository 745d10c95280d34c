#!/usr/bin/env python
"""Where does a bench.py step go?  torch.profiler kernel table + per-section CUDA-event timing of the
S3DIS tower (development aid; numbers quoted in profiles/ come from bench.py / bench_ops.py / ncu)."""
import sys
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


def step():
    opt.zero_grad(set_to_none=True)
    loss, _, _ = m(b["xyzs"], b["feats"], b["labels"], b["block_ids"], B)
    loss.backward()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
print(prof.key_averages(group_by_input_shape=True).table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=40))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=45, max_name_column_width=70))
# glue ops by input shape (what to fuse next)
rows = [e for e in prof.key_averages(group_by_input_shape=True)
        if e.key in ("aten::sum", "aten::cat", "aten::add", "aten::add_", "aten::mul", "aten::relu", "aten::threshold_backward",
                     "aten::native_dropout", "aten::native_dropout_backward", "aten::copy_", "aten::fill_", "aten::zero_",
                     "aten::mm", "aten::addmm", "aten::bmm", "aten::index", "aten::_log_softmax", "aten::nll_loss_forward",
                     "aten::slice_backward", "aten::split_with_sizes", "aten::clone", "aten::contiguous")]
rows.sort(key=lambda e: -e.self_device_time_total)
print("\n== glue by shape (self CUDA us over the 2 profiled steps)")
for e in rows[:70]:
    print(f"{e.key:32s} {e.self_device_time_total:9.0f} us  x{e.count:<4d} {str(e.input_shapes)[:110]}")

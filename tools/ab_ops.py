#!/usr/bin/env python
"""A/B of the kernel-dispatch knobs (pcs_set_option) on the op-level table of bench_ops.py.

    python tools/ab_ops.py --set seg_stream_max_row_bytes=0,walk_rows_narrow=4 --only sum_feat_gather --channels 3,35

Prints the same rows as bench_ops.py under the given option values (defaults otherwise)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

ap = argparse.ArgumentParser()
ap.add_argument("--set", default="")
ap.add_argument("--only", default="")
ap.add_argument("--channels", default="3,12,32,35,64,128,131")
ap.add_argument("--dtypes", default="f32")
ap.add_argument("--iters", type=int, default=7)
ap.add_argument("--no-ref", action="store_true")
a = ap.parse_args()

import bench_ops  # noqa: E402
from pointcloudsegmentation_b200 import ops  # noqa: E402

for kv in filter(None, a.set.split(",")):
    k, v = kv.split("=")
    ops.set_option(k, int(v))
print(f"## options: {a.set or '(defaults)'}", flush=True)
bench_ops.run_table(channels=[int(c) for c in a.channels.split(",")], dtypes=a.dtypes.split(","), iters=a.iters,
                    use_ref=not a.no_ref, only=set(filter(None, a.only.split(","))), log=lambda s: print(s, flush=True))

#!/usr/bin/env python
"""bench.py — S3DIS block points/s, forward+backward, on N B200s of one node.

One "step" = one training step of the S3DIS graph-pool tower (train_graph_pool_new.py:50-71:
points_pooling_two_layers -> pointnet_13_dilated_embed -> classifier_v3 -> weighted CE, backward,
gradient mean over ranks, Adam) on a block-diagonal batch of synthetic S3DIS-shaped blocks.  Every
custom op of the reference on that path (tf_ops/: 13 radius searches, voxel pooling, scatter /
gather / max-pool ops and their adjoints) runs in this repo's sm_100a kernels through the C ABI; the
fully_connected glue is stock PyTorch, as it is stock TensorFlow upstream.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--blocks B] [--impl reference]

N>1: launched by torchrun, one rank per GPU, NCCL; blocks are sharded across ranks (weak scaling: B
blocks per rank) and the only collective is the all-reduce of the flattened weight gradient.
Prints ONE JSON line on rank 0.  `--impl reference` times the CPU port of the same step (oracle/,
host cores) on a bounded sample — the reference itself has no CPU implementation of these ops.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "S3DIS block points/s fwd+bwd"
UNIT = "points/s"
WORKLOAD = ("configs[2]: S3DIS room blocks (3 m, stride-0.05 m surfaces, 6144-10000 pts/block), graph-pool "
            "segmentation tower pointnet_13_dilated_embed, train step fwd+bwd+Adam")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--blocks", type=int, default=16, help="blocks per GPU per step")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cpu-blocks", type=int, default=1, help="blocks per step of the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ops-report", default="", help="write the per-kernel table (JSON) to this path")
    return ap.parse_args()


# ---- algorithmic bytes per C-ABI call (SURVEY.md §8a "Alg. bytes"), from the call's own size args ----
def _alg_bytes(name, a):
    I = lambda k: int(a[k])  # noqa: E731
    if name == "pcs_neighbor_scatter":
        pn, c, cs = I(4), I(5), I(6); return 4 * c * (pn + cs) + 4 * cs + 8 * pn
    if name == "pcs_neighbor_gather":
        pn, c, cs = I(5), I(6), I(7); return 4 * c * (pn + cs) + 4 * cs + 8 * pn
    if name == "pcs_neighbor_sum_feat_gather":
        pn, c, cs = I(3), I(4), I(5); return 4 * c * (pn + cs) + 8 * pn
    if name == "pcs_neighbor_sum_feat_scatter":
        pn, c, cs = I(2), I(3), I(4); return 4 * c * (pn + cs) + 4 * cs
    if name == "pcs_neighbor_max_feat_gather":
        p1, p2, c = I(3), I(4), I(5); return 4 * c * p1 + 8 * c * p2 + 8 * p2
    if name == "pcs_neighbor_max_feat_scatter":
        p1, p2, c = I(4), I(5), I(6); return 8 * c * p2 + 4 * c * p1 + 8 * p2
    if name in ("pcs_location_weight_sum", "pcs_location_weight_sum_backward"):
        pn, m, cs = I(3), I(4), I(5); return 4 * m * (cs + pn) + 8 * pn
    if name == "pcs_location_weight_feat_sum":
        pn, m, o, cs = I(4), I(5), I(6), I(7); return 4 * cs * m * (o + 1) + 4 * pn * m * o + 8 * pn
    if name == "pcs_location_weight_feat_sum_backward":
        pn, m, o, cs = I(5), I(6), I(7), I(8); return 8 * cs * m * (o + 1) + 4 * pn * m * o + 8 * pn
    if name == "pcs_permutate_feature":
        pn, c = I(2), I(3); return 8 * c * pn + 4 * pn
    if name == "pcs_eliminate_center":
        pn, cs = I(3), I(4); return 4 * cs + 8 * pn + 8 * (cs - pn) + 8 * pn
    if name == "pcs_compute_voxel_index":
        return 24 * I(1)
    if name in ("pcs_compute_permutation_info", "pcs_compute_permutation_info64"):
        return 32 * I(3)
    if name == "pcs_compute_diff_xyz":
        return 28 * I(3) + 12 * I(4)
    if name == "pcs_compute_repermutation_info":
        return 12 * I(3) + 20 * I(4)
    if name == "pcs_csr_transpose":
        return 8 * I(2) + 4 * I(1)
    if name in ("pcs_search_neighborhood_count", "pcs_search_neighborhood_fill", "pcs_search_neighborhood_fixed"):
        return 20 * I(1)  # + 8*en for fill; the search is candidate-test bound, not byte bound
    return 0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}

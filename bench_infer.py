#!/usr/bin/env python
"""bench_infer.py — BASELINE.json configs[4]: Semantic3D-shaped dense block (~1M points) through the
inference-side op path: PoolingIndex two-level voxel pooling (0.45 / 1.5 m) + the multi-level neighbour
ops of model_pointnet_semantic3d.py:332-346 (radius 0.3, annuli 0.3-0.4 and 0.2-0.3 on level 0; 0.9 on
level 1), each graph followed by the ops an edge convolution issues in inference (diff-scatter of xyz,
concat-scatter of a 32-channel feature, max-gather).  Blocks are independent: with torchrun every rank
processes its own block, no communication; the line reports whole-job points/s.

    python bench_infer.py [--pn 1000000] [--steps 5] [--warmup 2]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pn", type=int, default=1_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--channels", type=int, default=32)
    args = ap.parse_args()
    import torch.distributed as dist

    from pointcloudsegmentation_b200 import launch_count
    from pointcloudsegmentation_b200 import ops as P
    from pointcloudsegmentation_b200.graph_layers import (graph_concat_scatter, graph_pool, points_pooling_two_layers,
                                                          search_neighborhood, search_neighborhood_range)
    from pointcloudsegmentation_b200.synthetic import semantic3d_block

    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    x = torch.from_numpy(semantic3d_block(rank, args.pn)).cuda()
    f = torch.rand(args.pn, 4, device="cuda")
    lab = torch.zeros(args.pn, dtype=torch.int64, device="cuda")
    C = args.channels
    stages = {}

    def tick(name, e0):
        e1 = torch.cuda.Event(enable_timing=True); e1.record()
        stages.setdefault(name, []).append((e0, e1))
        e2 = torch.cuda.Event(enable_timing=True); e2.record()
        return e2

    def step():
        e = torch.cuda.Event(enable_timing=True); e.record()
        xs, dxs, ff, _, vl, vb, vc = points_pooling_two_layers(x, f, lab, 0.45, 1.5, 80.0)
        e = tick("pooling (2 levels)", e)
        edges = 0
        feats = [torch.ones(xs[0].shape[0], C, device="cuda"), torch.ones(xs[1].shape[0], C, device="cuda")]
        for lvl, rmin, rmax in ((0, None, 0.3), (0, 0.3, 0.4), (0, 0.2, 0.3), (1, None, 0.9)):
            g = (search_neighborhood(xs[lvl], rmax) if rmin is None else search_neighborhood_range(xs[lvl], rmin, rmax))
            e = tick(f"search L{lvl} {rmin or 0}-{rmax}", e)
            nidxs, nlens, nbegs, ncens = g
            edges += int(nidxs.shape[0])
            sx = P.neighbor_scatter(xs[lvl], nidxs, nlens, nbegs, True)
            sf = graph_concat_scatter(feats[lvl], nidxs, nlens, nbegs, ncens)
            out = graph_pool(sf[:, :C].contiguous(), nlens, nbegs)
            e = tick(f"edge ops L{lvl} {rmin or 0}-{rmax}", e)
            del sx, sf, out
        return edges

    with torch.no_grad():
        for _ in range(args.warmup):
            edges = step()
        stages.clear()
        torch.cuda.synchronize()
        l0 = launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            edges = step()
        e1.record()
        torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms) / args.steps
    if rank == 0:
        out = {"metric": "Semantic3D dense block points/s, inference op path (pooling + 4 graphs + edge ops)",
               "value": world * args.pn / (ms * 1e-3), "unit": "points/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "dtype": "f32",
               "data": "synthetic", "config": {"workload": "configs[4]: Semantic3D dense block", "points_per_block": args.pn,
                                               "edges_per_block": edges, "channels": C, "blocks_per_gpu": 1},
               "gpu_launches": launch_count() - l0,
               "stages_ms": {k: round(float(np.mean([a.elapsed_time(b) for a, b in v])), 3) for k, v in stages.items()}}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

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
# --dataset scannet = BASELINE configs[3]: the same tower on ScanNet-shaped input (train_gpn_scannet.py:44-65:
# rgb + covariance features, 21 classes); blocks come from the same surface generator.
DATASETS = {"s3dis": dict(feat_dim=3, num_classes=13, workload=WORKLOAD),
            "scannet": dict(feat_dim=12, num_classes=21,
                            workload="configs[3]: ScanNet-shaped blocks (rgb + 9 covariance features, 21 classes), "
                                     "graph-pool segmentation tower pointnet_13_dilated_embed, train step fwd+bwd+Adam, "
                                     "data-parallel with one NCCL gradient all-reduce"),
            # the same S3DIS blocks through model_pgnet's edge-conditioned-diffusion tower (unfused edge MLPs)
            "s3dis_pgnet": dict(feat_dim=3, num_classes=13, model="pgnet", default_blocks=32,
                                workload="configs[2] with model_pgnet: S3DIS room blocks, pgnet_model_v3 (three ecd_stage: "
                                         "ecd_xyz + 2 / 9 / 9 ecd_feats layers) + classifier_v3, train step fwd+bwd+Adam; custom ops "
                                         "in this repo's kernels, edge MLPs stock PyTorch on materialised edge tensors"),
            # BASELINE configs[1]: the anchor-weight classification network (run_modelnet below)
            "modelnet40": dict(feat_dim=9, num_classes=40,
                               workload="configs[1]: ModelNet40-shaped clouds (1024 points on box / cylinder surfaces in the unit "
                                        "sphere), batch 32 per GPU, train step fwd+bwd+Adam of graph_conv_pool_model_v1 + "
                                        "model_classifier_v1 (train_modelnet40.py:41-58; anchor-weight convolutions on the "
                                        "location-weight ops), hierarchy (covariances, 2 voxel levels, 3 radius graphs) built on "
                                        "the GPU inside the step"),
            # BASELINE configs[4]: inference, tiles sharded over the ranks, no collective (run_semantic3d below)
            "semantic3d": dict(feat_dim=4, num_classes=8,
                               workload="configs[4]: Semantic3D dense tiles (80 m x 80 m = 64 of the reference's 10 m blocks, "
                                        "~1M points at ~0.1 m), INFERENCE of the dense model (train_gpn_semantic3d_dense.py:49-85: "
                                        "dense_feats over r=0.25 neighbourhoods of the 0.15 m grid centres -> "
                                        "points_pooling_two_layers(0.45, 1.5, 10.0) -> pointnet_13_dilate_embed_semantic3d -> "
                                        "classifier_v3), block preparation (voxel sort, grid down-sample, dense radius search) "
                                        "inside the step; tiles sharded over the ranks, no collective")}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--blocks", type=int, default=0,
                    help="blocks per GPU per step (default 64; 32 for s3dis_pgnet, whose unfused edge tensors cost ~1 GB per block; the step is "
                         "host-launch-bound below ~48: 16 blocks = 24 ms of enqueueing for 17 ms of GPU work)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dataset", default="s3dis", choices=sorted(DATASETS), help="default = the headline config")
    ap.add_argument("--cpu-blocks", type=int, default=1, help="blocks per step of the CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ops-report", default="", help="write the per-kernel table (JSON) to this path")
    ap.add_argument("--tile-points", type=int, default=1_000_000, help="--dataset semantic3d: dense points per tile")
    ap.add_argument("--no-other-configs", action="store_true",
                    help="skip the short runs of the other BASELINE configs (`other_configs` key, N=1 headline run only)")
    ap.add_argument("--no-ops", action="store_true", help="skip the per-op roofline table (`ops` key, N=1 only)")
    ap.add_argument("--no-reference-cuda", action="store_true",
                    help="skip the `reference_cuda` block (the tower on the reference's own CUDA kernels, N=1 only)")
    return ap.parse_args()


# ---- algorithmic bytes per C-ABI call (SURVEY.md §8a "Alg. bytes"), from the call's own size args ----
def _alg_bytes(name, a):
    I = lambda k: int(a[k])  # noqa: E731
    if name == "pcs_neighbor_scatter":
        pn, c, cs = I(5), I(6), I(7); return 4 * c * (pn + cs) + 4 * cs + 8 * pn
    if name == "pcs_neighbor_gather":
        pn, c, cs = I(5), I(6), I(7); return 4 * c * (pn + cs) + 4 * cs + 8 * pn
    if name == "pcs_neighbor_sum_feat_gather":
        pn, c, cs = I(3), I(4), I(5); return 4 * c * (pn + cs) + 8 * pn
    if name == "pcs_neighbor_sum_feat_scatter":
        pn, c, cs = I(4), I(5), I(6); return 4 * c * (pn + cs) + 4 * cs
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
    # fused edge-conv extension (csrc/dense_mlp.cu, csrc/edge_conv.cu): compulsory reads + writes per row
    if name == "pcs_dense_mlp3_forward":
        rows, A, B, C, O = (I(k) for k in range(4, 9)); return 4 * rows * (A + B + C + 2 * O)
    if name == "pcs_dense_mlp3_backward":
        rows, A, B, C, O = (I(k) for k in range(5, 10)); return 4 * rows * (3 * (A + B + C) + 3 * O)
    if name == "pcs_dense_mlp3_backward_pooled":
        rows, A, B, C, O = (I(k) for k in range(8, 13)); return 4 * rows * (3 * (A + B + C) + 2 * O + 1)
    if name == "pcs_pool_linear_wgrad":
        pn, cs, O, J = I(5), I(6), I(7), I(8); return 8 * pn * O + 8 * pn + 4 * J * min(cs, pn * O)
    if name == "pcs_edge_combine_forward":
        pn, cs, H = I(8), I(9), I(10); return cs * (4 * H + 16) + pn * (8 * H + 8)
    if name == "pcs_edge_xyz_wgrad":
        cs, H = I(2), I(3); return cs * (4 * H + 12)
    if name == "pcs_column_sum":
        rows, C = I(1), I(2); return 4 * C * (rows + 1)
    if name == "pcs_linear_tc_wgrad":  # g and x read once, the partials written once
        M, N, K = I(1), I(2), I(5); return 4 * (M * N + M * K + N * (K + 1))
    if name == "pcs_linear_tc":  # x read once, y written once, the weight once, the addend when there is one
        M, K, N = I(1), I(2), I(5); return 4 * (M * K + M * N * (2 if a[9] else 1) + N * K)
    if name == "pcs_relu_concat_dropout_forward":  # a, b read once, out written once
        M, Na, Nb = I(1), I(2), I(5); return 8 * M * (Na + Nb)
    if name == "pcs_relu_concat_dropout_backward":  # g read, out[:, :Na] read, da and db written
        M, Na, Nb = I(2), I(3), I(4); return 4 * M * (3 * Na + 2 * Nb)
    return 0


def _alg_flops(name, a):
    """fp32 FMA work (2 flop per MAC) of the FLOP-bound fused edge-MLP kernels; 0 for the byte movers."""
    I = lambda k: int(a[k])  # noqa: E731
    if name == "pcs_dense_mlp3_forward":
        rows, A, B, C, O = (I(k) for k in range(4, 9)); return 2 * rows * (B * A + C * (B + A) + O * (A + B + C))
    if name in ("pcs_dense_mlp3_backward", "pcs_dense_mlp3_backward_pooled"):
        rows, A, B, C, O = (I(k) for k in (range(5, 10) if name.endswith("backward") else range(8, 13)))
        return 2 * rows * (2 * (B * A + C * (B + A)) + O * (A + B + C))  # recompute h + the three dX products
    if name == "pcs_linear_tc":  # fp32 products of the layer (each is issued as 3 TF32 MMAs)
        return 2 * I(1) * I(2) * I(5)
    if name == "pcs_linear_tc_wgrad":
        return 2 * I(1) * I(2) * I(5)
    return 0


# fp32 FFMA peak of one B200: 148 SMs x 128 lanes x 2 flop x SM clock (nominal, from the sampled clock)
def fp32_peak_tflops(sm_mhz):
    return 148 * 128 * 2 * sm_mhz * 1e6 / 1e12


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines, self.first = index, None, [], 0

    def wait_first(self, timeout=5.0):
        t0 = time.time()
        while self.proc is not None and not self.lines and time.time() - t0 < timeout:
            time.sleep(0.05)

    def mark(self):
        """Samples from here on belong to the timed region (the sampler is started before the warm-up:
        nvidia-smi's own start-up initialises NVML, which briefly holds driver locks the enqueueing thread needs)."""
        self.first = len(self.lines)

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
        for ln in self.lines[self.first:]:
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


# ---- CPU arm: the oracle port of the same step on the host cores ------------------------------------
def tower_class(dataset):
    """The tower of a train-step dataset: pointnet_13_dilated_embed, or model_pgnet's pgnet_model_v3."""
    if DATASETS[dataset].get("model") == "pgnet":
        from pointcloudsegmentation_b200.pgnet import PGNetV3

        return PGNetV3
    from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed

    return Pointnet13DilatedEmbed


def cpu_step_fn(blocks_per_step, seed=12345, dataset="s3dis"):
    """Builds the CPU network (oracle-backed layer namespace) and returns (step(), points_per_step)."""
    import torch

    from oracle import cpu_layers
    from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed
    from pointcloudsegmentation_b200.synthetic import s3dis_batch

    torch.manual_seed(0)
    ds = DATASETS[dataset]
    data = s3dis_batch(seed, blocks_per_step, feat_dim=ds["feat_dim"], num_classes=ds["num_classes"])
    x, f, l, b = (torch.from_numpy(data[k]) for k in ("xyzs", "feats", "labels", "block_ids"))
    # the reference's literal evaluation order
    if ds.get("model") == "pgnet":
        from oracle import torch_ops  # noqa: F401

        model = tower_class(dataset)(ds["num_classes"], L=cpu_layers.layers(), ops=sys.modules["oracle.torch_ops"])
    else:
        model = Pointnet13DilatedEmbed(ds["num_classes"], L=cpu_layers.layers(), split=False,
                                       class_weights=[1.0] * ds["num_classes"] if dataset != "s3dis" else None)
    model.train()
    with torch.no_grad():
        model(x, f, l, b, blocks_per_step)  # materialise the lazy layers
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)

    def step():
        opt.zero_grad(set_to_none=True)
        loss, _, _ = model(x, f, l, b, blocks_per_step)
        loss.backward()
        opt.step()
        return float(loss.detach())

    return step, int(x.shape[0])


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every core this process may run on."""
    import torch

    from oracle import pcs_oracle

    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(n)
    pcs_oracle.set_num_threads(n)
    return n


def cpu_threads():
    import torch

    from oracle import pcs_oracle

    return max(pcs_oracle.num_threads(), torch.get_num_threads())


def run_reference(args):
    """`--impl reference`: the CPU port, all host threads, bounded sample (cpu_blocks blocks/step)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    use_all_host_threads()
    sem = args.dataset == "semantic3d"
    step, pts = (semantic3d_cpu_step() if sem else modelnet_cpu_step() if args.dataset == "modelnet40"
                 else cpu_step_fn(args.cpu_blocks, dataset=args.dataset))
    for _ in range(max(args.warmup, 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    v = pts * args.steps / dt
    sample = (f"4 clouds ({pts} points) per step, {args.steps} steps" if args.dataset == "modelnet40" else
              f"one 10 m x 10 m crop of a tile ({pts} dense points) per step, {args.steps} steps" if sem else
              f"{args.cpu_blocks} block(s)/step ({pts} points) of the same generator, {args.steps} steps")
    metric = {"s3dis": METRIC, "s3dis_pgnet": METRIC + " (model_pgnet tower)", "scannet": METRIC.replace("S3DIS", "ScanNet-shaped"),
              "semantic3d": "Semantic3D dense tile points/s inference", "modelnet40": "ModelNet40 points/s fwd+bwd"}[args.dataset]
    out = {"metric": metric, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": max(args.warmup, 1), "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
           "config": {"workload": DATASETS[args.dataset]["workload"], "blocks_per_step": args.cpu_blocks, "points_per_step": pts},
           "cpu_baseline": {"value": v, "unit": UNIT, "cores": cpu_threads(), "kind": "port", "sample": sample},
           "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(out), flush=True)
    return 0


def pin_rank_threads(local, world):
    """N ranks x (enqueueing thread + look-ahead thread + autograd thread) on one host: give every rank its own slice
    of the cores this job may run on, so that the ranks' Python threads do not migrate over each other."""
    if world <= 1 or not hasattr(os, "sched_setaffinity") or os.environ.get("PCSEG_NO_PIN"):
        return None
    cores = sorted(os.sched_getaffinity(0))
    per = len(cores) // world
    if per < 2:
        return None
    mine = cores[local * per:(local + 1) * per]
    try:
        os.sched_setaffinity(0, mine)
    except OSError:
        return None
    return len(mine)


# ---- GPU arm ---------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import pointcloudsegmentation_b200 as pcs
    from pointcloudsegmentation_b200._lib import profile
    from pointcloudsegmentation_b200.dist import GradAverager, broadcast_parameters
    from pointcloudsegmentation_b200.pipeline import PrepareAhead
    from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed
    from pointcloudsegmentation_b200.synthetic import s3dis_batch_to_budget

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    pin_rank_threads(local, world)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL's banner / debug output goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    ds = DATASETS[args.dataset]
    K, W, B = args.steps, max(args.warmup, 3), (args.blocks or ds.get("default_blocks", 64))

    # distinct batches per rank (weak scaling: B blocks per rank per step), host copies pinned
    n_distinct = 4
    host, resident = [], []
    for i in range(n_distinct):
        # every rank's loader fills its step to the same point budget (B blocks out of B + 8 candidates, mean block
        # size x B): a synchronous step is as slow as its heaviest rank
        d = s3dis_batch_to_budget(1000 * rank + i, B, B * (6144 + 10000) // 2, feat_dim=ds["feat_dim"], num_classes=ds["num_classes"])
        h = {k: torch.from_numpy(d[k]).pin_memory() for k in ("xyzs", "feats", "labels", "block_ids")}
        host.append(h)
        resident.append({k: v.to(dev) for k, v in h.items()})
    pts_per_step = [int(h["xyzs"].shape[0]) for h in host]
    h2d_bytes = int(np.mean([sum(v.numel() * v.element_size() for v in h.values()) for h in host]))

    torch.manual_seed(0)
    if ds.get("model") == "pgnet":
        model = tower_class(args.dataset)(ds["num_classes"]).to(dev)
    else:
        model = Pointnet13DilatedEmbed(ds["num_classes"],
                                       class_weights=[1.0] * ds["num_classes"] if args.dataset != "s3dis" else None).to(dev)
    model.train()
    with torch.no_grad():
        r0 = resident[0]
        model(r0["xyzs"], r0["feats"], r0["labels"], r0["block_ids"], B)  # materialise lazy layers
    torch.cuda.synchronize()
    params = [p for p in model.parameters()]
    broadcast_parameters(params)
    opt = torch.optim.Adam(params, lr=1e-3, fused=True)
    n_param = sum(p.numel() for p in params)
    average_grads = GradAverager(params)

    # The point-only part of a step (voxel pooling + the nine radius searches, model.prepare) runs one
    # batch ahead on a second stream from a second host thread, like the reference's provider threads: its
    # host syncs (edge counts come back to size the index tensors) then stall neither the GPU nor the
    # enqueueing of the training step.  Every timed step still contains exactly one prepare (of the next
    # batch) and one fwd+bwd+Adam.
    main = torch.cuda.current_stream()
    ahead = PrepareAhead(model, B, dev)  # own stream + own host thread (pipeline.py)
    side = ahead.stream
    pending = {}

    def prepare_ahead(key, batch_fn):
        pending.clear()
        pending[key] = ahead.submit(batch_fn)

    def take(key, batch_fn):
        if key not in pending:
            prepare_ahead(key, batch_fn)
        return ahead.take(pending.pop(key))

    def join_ahead():
        for f in pending.values():  # the look-ahead job must have been fully enqueued before its stream is waited on
            f.result()

    ar_events = []

    def train_step(prep):
        opt.zero_grad(set_to_none=True)
        loss, _, _ = model(prepared=prep)
        loss.backward()
        if world > 1:  # device time of the collective (pack + all-reduce + unpack), two event records per step
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            average_grads()  # the only collective on the path: mean of weight gradients (train_util.py:77-111)
            a1.record()
            ar_events.append((a0, a1))
        else:
            average_grads()
        opt.step()
        return loss

    def resident_batch(i):
        return lambda: resident[i % n_distinct]

    def host_batch(i):
        h = host[i % n_distinct]
        return lambda: {k: v.to(dev, non_blocking=True) for k, v in h.items()}

    def step_resident(i):
        loss = train_step(take(("r", i), resident_batch(i)))
        prepare_ahead(("r", i + 1), resident_batch(i + 1))
        return loss

    # e2e: every step's loss is copied to pinned host memory and READ on the host; the read of step i is
    # deferred until step i+1 has been enqueued (the way a training loop logs without stalling the queue),
    # and the last one is drained inside the timed region.
    losses, inflight = [], []

    def read_loss():
        buf, ev = inflight.pop(0)
        ev.synchronize()
        losses.append(float(buf))

    def step_e2e(i):
        loss = train_step(take(("h", i), host_batch(i)))
        buf = torch.empty((), dtype=torch.float32).pin_memory()
        buf.copy_(loss.detach(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        inflight.append((buf, ev))
        prepare_ahead(("h", i + 1), host_batch(i + 1))  # includes the next batch's H2D copy
        while len(inflight) > 1:
            read_loss()

    def drain_e2e():
        while inflight:
            read_loss()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = []

    def timed(fn, n, offset=0, finish=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for i in range(n):
            fn(offset + i)
        if finish:
            finish()
        join_ahead()
        main.wait_stream(side)  # the look-ahead prepare issued by the last step belongs to the timed region
        e1.record()
        host_ms.append(1e3 * (time.perf_counter() - t0))  # time the host needed to enqueue the region
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    # shape priming (untimed, before the W warm-up steps): the n_distinct synthetic batches have different
    # point / edge counts, so each is stepped once first — cuBLAS picks its algorithm per GEMM shape and the
    # caching allocator grows on first sight of a size; neither belongs to a steady-state step
    for i in range(n_distinct):
        step_resident(i)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()  # NVML is up before the warm-up steps begin
    for i in range(W):
        step_resident(i)
    torch.cuda.reset_peak_memory_stats()
    sampler.mark()
    profile.enabled = False
    l0 = pcs.launch_count()
    ar_events.clear()
    ms = timed(step_resident, K, W)
    launches = pcs.launch_count() - l0
    allreduce_us = None
    if world > 1:
        ar = torch.tensor([1e3 * float(np.mean([a.elapsed_time(b) for a, b in ar_events]))], device=dev)
        dist.all_reduce(ar, op=dist.ReduceOp.MAX)
        allreduce_us = float(ar)
    peak_mem = torch.cuda.max_memory_allocated()
    # second timed region over the same K steps, now with a CUDA event pair around every launch of this
    # library (per-kernel attribution + the roofline's launch durations).  Kept out of the headline region:
    # ~9000 event records per step cost host time, and the step is within 10 % of being launch-bound.
    profile.records.clear()
    profile.enabled = True
    ms_attr = timed(step_resident, K, W)
    profile.enabled = False
    clocks = sampler.stop() if rank == 0 else None
    pts = torch.tensor([float(sum(pts_per_step[(W + i) % n_distinct] for i in range(K)))], device=dev)
    if world > 1:
        dist.all_reduce(pts)
    value = float(pts) / (ms * 1e-3)

    # per-kernel attribution from the events recorded inside the timed region
    table = {}
    for name, a, e0, e1 in profile.records:
        t = table.setdefault(name, {"calls": 0, "ms": 0.0, "bytes": 0, "flops": 0})
        t["calls"] += 1
        t["ms"] += e0.elapsed_time(e1)
        t["bytes"] += _alg_bytes(name, a)
        t["flops"] += _alg_flops(name, a)
    profile.records.clear()

    step_e2e(0)
    drain_e2e()
    ms_e2e = timed(step_e2e, K, 1, finish=drain_e2e)
    assert len(losses) == K + 1 and all(np.isfinite(losses)), "e2e: every step's loss must have been read"
    pts2 = torch.tensor([float(sum(pts_per_step[(1 + i) % n_distinct] for i in range(K)))], device=dev)
    if world > 1:
        dist.all_reduce(pts2)
    e2e_value = float(pts2) / (ms_e2e * 1e-3)
    join_ahead()
    ahead.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return dict(world=world, rank=rank, K=K, W=W, B=B, ms=ms, ms_attr=ms_attr, host_ms=host_ms[0], value=value, e2e=e2e_value, table=table,
                launches=launches, clocks=clocks, peak_mem=peak_mem, h2d=h2d_bytes, n_param=n_param, allreduce_us=allreduce_us,
                pts_per_step=int(np.mean(pts_per_step)))


# ---- N=1 extras of the driver-run line: single-block steps (the reference's unit of work) on this repo's kernels and
# ---- on the reference's own CUDA kernels, and the per-op roofline table -----------------------------------------------
def single_block_steps(dataset, impl, steps, warmup=2, n_distinct=4):
    """fwd+bwd+Adam on ONE block per step with prepare (voxel pooling + searches) inline on the same stream — how the
    reference feeds a tower (train_graph_pool.py:95-108,143-159).  impl = "b200" (this repo's kernels, split/fused
    evaluation) or "reference_cuda" (the reference's unmodified tf_ops kernels through oracle/_ref/libref.so, literal
    evaluation order, stock PyTorch MLPs).  -> dict(ms_per_step, points_per_step, value) or None if unavailable."""
    import torch

    from pointcloudsegmentation_b200.network import Pointnet13DilatedEmbed
    from pointcloudsegmentation_b200.synthetic import s3dis_batch

    ds = DATASETS[dataset]
    kw = dict(class_weights=[1.0] * ds["num_classes"] if dataset != "s3dis" else None)
    if impl == "reference_cuda":
        from oracle import libref, libref_layers

        if not libref.available():
            return None
        libref.SYNC = False  # stream order only: the reference launches on the legacy default stream = torch's default
        kw.update(L=libref_layers.layers(), split=False)
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(0)
    model = Pointnet13DilatedEmbed(ds["num_classes"], **kw).to(dev)
    model.train()
    batches = []
    for i in range(n_distinct):
        d = s3dis_batch(5000 + i, 1, feat_dim=ds["feat_dim"], num_classes=ds["num_classes"])
        batches.append(tuple(torch.from_numpy(d[k]).to(dev) for k in ("xyzs", "feats", "labels")))
    with torch.no_grad():
        model(*batches[0])
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)

    def step(i):
        x, f, l = batches[i % n_distinct]
        opt.zero_grad(set_to_none=True)
        loss, _, _ = model(x, f, l)
        loss.backward()
        opt.step()
        return loss

    for i in range(max(warmup, n_distinct)):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        loss = step(i)
    e1.record()
    torch.cuda.synchronize()
    assert bool(torch.isfinite(loss)), "single-block step produced a non-finite loss"
    ms = e0.elapsed_time(e1) / steps
    pts = float(np.mean([batches[i % n_distinct][0].shape[0] for i in range(steps)]))
    if impl == "reference_cuda":
        libref.SYNC = True
    return {"ms_per_step": ms, "points_per_step": pts, "value": pts / (ms * 1e-3), "unit": UNIT, "steps": steps,
            "blocks_per_step": 1}


def ops_table():
    """The other half of BASELINE.json's metric: per-op achieved HBM GB/s vs roofline on configs[0], with the reference's
    own CUDA kernel timed beside each f32 row (bench_ops.run_table).  Rows are compact lists, see `columns`."""
    import bench_ops

    t = bench_ops.run_table(iters=5, log=lambda *a: None)
    cols = ["op", "dtype", "C", "blocks", "us", "GBps", "frac", "ref_cuda_us"]
    rnd = lambda v: round(v, 4 if isinstance(v, float) and v < 10 else 1) if isinstance(v, float) else v  # noqa: E731
    return {"workload": "configs[0]: 4096-pt S3DIS-shaped blocks, radius-0.1 CSR (mean degree 12), block-diagonal batch "
                        f"sized so that every row moves >= {t['min_alg_mb']:.0f} MB (>> L2); L2 flushed between iterations",
            "hbm_peak_gbs": t["hbm_peak_gbs"], "hbm_peak_source": t["hbm_peak_source"], "iters": t["iters"],
            "min_frac_over_ops": t["min_frac_over_ops"], "ops_below_70pct": t["ops_below_70pct"],
            "n_rows": len(t["rows"]), "columns": cols, "rows": [[rnd(r[c]) for c in cols] for r in t["rows"]]}


# ---- BASELINE configs[4]: Semantic3D dense-tile inference, tiles sharded over the ranks -------------------------------
def semantic3d_tile(seed, pn):
    """One synthetic dense tile, block-normalised as semantic3d_dense_util.normalize does (min corner at the origin):
    xyz [pn,3], feats [pn,4] (rgb + intensity), labels [pn] in 0..8 (0 = unlabelled)."""
    from pointcloudsegmentation_b200.synthetic import semantic3d_block

    rng = np.random.default_rng(10_000 + seed)
    x = semantic3d_block(seed, pn)
    x -= x.min(axis=0, keepdims=True)
    return {"xyzs": x, "feats": rng.random((pn, 4), dtype=np.float32), "labels": rng.integers(0, 9, pn).astype(np.int64)}


def semantic3d_prepare_fn(model, P, PU):
    """-> prepare(batch) -> everything of a step that reads no parameter: block preparation of the dense tile (voxel
    sort at the centre stride, grid down-sample, dense r = 0.25 radius search) + the centres' voxel pooling and graphs."""

    def prepare(b):
        x, f, l = b["xyzs"], b["feats"], b["labels"]
        # voxel sort of the dense cloud at the centre stride: the first point of every voxel is that voxel's centre
        # (gridDownsampleGPU(xyzs, 0.15), train_gpn_semantic3d_dense.py:365) and the sorted order is what makes the
        # culled searches effective
        o2p, _, _, begs, _ = P.compute_permutation_info(PU._voxel_index_from_min(x, 0.15))
        xs, fs = P.permutate_feature(x, o2p), P.permutate_feature(f, o2p)
        ls = l[o2p.long()]
        csr = PU.findNeighborRadiusCPU(xs, begs, 0.25)  # semantic3d_dense_util.py:8,33
        return xs, fs, ls, begs, csr, model.prepare_points(xs, ls, begs)

    return prepare


def semantic3d_forward(model, prep):
    import torch

    xs, fs, ls, begs, (nidxs, nlens, nbegs, ncens), points = prep
    with torch.no_grad():
        _, logits, _, _ = model(xs, fs, ls, begs, nidxs, nlens, nbegs, ncens, points=points)
    return logits, int(begs.shape[0])


def run_semantic3d(args):
    import torch
    import torch.distributed as dist

    import pointcloudsegmentation_b200 as pcs
    from pointcloudsegmentation_b200 import ops as P
    from pointcloudsegmentation_b200 import point_util as PU
    from pointcloudsegmentation_b200._lib import profile
    from pointcloudsegmentation_b200.network import Semantic3DDense

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    pin_rank_threads(local, world)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)  # only for the barrier and the max over ranks of the timing
    K, W, N = args.steps, max(args.warmup, 3), args.tile_points
    n_distinct = 2
    host = [{k: torch.from_numpy(v).pin_memory() for k, v in semantic3d_tile(100 * rank + i, N).items()} for i in range(n_distinct)]
    resident = [{k: v.to(dev) for k, v in h.items()} for h in host]
    h2d = int(sum(v.numel() * v.element_size() for v in host[0].values()))
    torch.manual_seed(0)
    model = Semantic3DDense().to(dev).eval()
    prepare = semantic3d_prepare_fn(model, P, PU)
    small = {k: v[:100_000].contiguous() for k, v in resident[0].items()}
    semantic3d_forward(model, prepare(small))  # materialise the lazy layers
    n_param = sum(p.numel() for p in model.parameters())
    main = torch.cuda.current_stream()
    # block preparation of tile i+1 (it contains the host syncs that size the index tensors) runs on a second stream
    # from a second host thread while the model runs on tile i; every timed step = 1 preparation + 1 forward
    from pointcloudsegmentation_b200.pipeline import PrepareAhead

    ahead = PrepareAhead(model, 1, dev, prepare_fn=prepare)
    side = ahead.stream
    pending = {}
    centres = []

    def prepare_ahead(key, batch_fn):
        pending.clear()
        pending[key] = ahead.submit(batch_fn)

    def take(key, batch_fn):
        if key not in pending:
            prepare_ahead(key, batch_fn)
        return ahead.take(pending.pop(key))

    def join_ahead():
        for f in pending.values():
            f.result()

    def resident_batch(i):
        return lambda: resident[i % n_distinct]

    def host_batch(i):
        h = host[i % n_distinct]
        return lambda: {k: v.to(dev, non_blocking=True) for k, v in h.items()}

    def step_resident(i):
        logits, m = semantic3d_forward(model, take(("r", i), resident_batch(i)))
        prepare_ahead(("r", i + 1), resident_batch(i + 1))
        centres.append(m)

    d2h = [0]
    inflight = []

    def read_logits():
        out, ev = inflight.pop(0)
        ev.synchronize()
        assert bool(np.isfinite(out.numpy()).all())

    def step_e2e(i):
        logits, m = semantic3d_forward(model, take(("h", i), host_batch(i)))
        out = torch.empty(logits.shape, dtype=torch.float32).pin_memory()  # the per-centre logits are the result
        out.copy_(logits, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        inflight.append((out, ev))
        d2h[0] = out.numel() * 4
        prepare_ahead(("h", i + 1), host_batch(i + 1))  # includes the next tile's H2D copy
        while len(inflight) > 1:  # the read of tile i is deferred until tile i+1 has been enqueued
            read_logits()

    def drain_e2e():
        while inflight:
            read_logits()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = []

    def timed(fn, n, offset=0, finish=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for i in range(n):
            fn(offset + i)
        if finish:
            finish()
        join_ahead()
        main.wait_stream(side)  # the look-ahead preparation issued by the last step belongs to the timed region
        e1.record()
        host_ms.append(1e3 * (time.perf_counter() - t0))
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    for i in range(W):
        step_resident(i)
    torch.cuda.reset_peak_memory_stats()
    sampler.mark()
    profile.enabled = False
    l0 = pcs.launch_count()
    centres.clear()
    ms = timed(step_resident, K, W)
    launches = pcs.launch_count() - l0
    peak_mem = torch.cuda.max_memory_allocated()
    profile.records.clear()
    profile.enabled = True
    ms_attr = timed(step_resident, K, W)
    profile.enabled = False
    clocks = sampler.stop() if rank == 0 else None
    table = {}
    for name, a, e0, e1 in profile.records:
        tt = table.setdefault(name, {"calls": 0, "ms": 0.0, "bytes": 0, "flops": 0})
        tt["calls"] += 1
        tt["ms"] += e0.elapsed_time(e1)
        tt["bytes"] += _alg_bytes(name, a)
        tt["flops"] += _alg_flops(name, a)
    profile.records.clear()
    step_e2e(0)
    drain_e2e()
    ms_e2e = timed(step_e2e, K, 1, finish=drain_e2e)
    join_ahead()
    ahead.close()
    pts = torch.tensor([float(N * K)], device=dev)
    if world > 1:
        dist.all_reduce(pts)
        dist.barrier()
        dist.destroy_process_group()
    return dict(world=world, rank=rank, K=K, W=W, B=1, ms=ms, ms_attr=ms_attr, host_ms=host_ms[0], value=float(pts) / (ms * 1e-3),
                e2e=float(pts) / (ms_e2e * 1e-3), table=table, launches=launches, clocks=clocks, peak_mem=peak_mem, h2d=h2d,
                d2h=d2h[0], n_param=n_param, pts_per_step=N, allreduce_us=None,
                extra_config={"centres_per_tile": int(np.mean(centres)), "tiles_per_gpu_per_step": 1,
                              "parallelism": f"{world} rank(s), tiles sharded, no collective",
                              "pipeline": "block preparation (voxel sort, grid down-sample, dense radius search, pooling indices, "
                                          "9 graphs) of tile i+1 runs on a second stream from a second host thread while the "
                                          "model runs on tile i; each timed step = 1 preparation + 1 forward"})


def semantic3d_cpu_step():
    """The CPU port of the same inference step (oracle layers) on a bounded sample: one 10 m x 10 m crop of a tile =
    one of the reference's blocks.  -> (step(), dense points per step)."""
    import torch

    from oracle import cpu_layers, point_util_oracle as O
    from pointcloudsegmentation_b200.network import Semantic3DDense

    t = semantic3d_tile(0, 1_000_000)
    m = (t["xyzs"][:, 0] >= 30) & (t["xyzs"][:, 0] < 40) & (t["xyzs"][:, 1] >= 30) & (t["xyzs"][:, 1] < 40)
    x = t["xyzs"][m] - t["xyzs"][m].min(axis=0, keepdims=True)
    f, l = t["feats"][m], t["labels"][m]
    torch.manual_seed(0)
    model = Semantic3DDense(L=cpu_layers.layers(), split=False).eval()

    def step():
        v = np.floor((x - x.min(0)) / np.float32(0.15)).astype(np.int64)
        o2p = np.argsort((v[:, 0] << 32) + (v[:, 1] << 16) + v[:, 2], kind="stable")
        xs, fs, ls = x[o2p], f[o2p], l[o2p]
        idxs = O.grid_downsample(xs, 0.15)
        nidxs, nlens, nbegs, ncens = O.cross_search(xs, xs[idxs], 0.25 * 0.25)
        with torch.no_grad():
            _, logits, _, _ = model(*(torch.from_numpy(np.ascontiguousarray(a)) for a in (xs, fs, ls, idxs, nidxs, nlens, nbegs, ncens)))
        return float(logits.sum())

    return step, int(x.shape[0])


# ---- BASELINE configs[1]: ModelNet40-shaped classification, train step ------------------------------------------------
def run_modelnet(args):
    import torch
    import torch.distributed as dist

    import pointcloudsegmentation_b200 as pcs
    from pointcloudsegmentation_b200._lib import profile
    from pointcloudsegmentation_b200.dist import GradAverager, broadcast_parameters
    from pointcloudsegmentation_b200.modelnet import GraphConvPoolModelV1
    from pointcloudsegmentation_b200.synthetic import modelnet_batch

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback)"
    pin_rank_threads(local, world)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    K, W, B, PN = args.steps, max(args.warmup, 3), 32, 1024
    n_distinct = 4
    host = []
    for i in range(n_distinct):
        d = modelnet_batch(1000 * rank + i, B, PN)
        rng = np.random.default_rng(77 + 1000 * rank + i)
        host.append({"xyzs": torch.from_numpy(d["xyzs"]).pin_memory(), "block_ids": torch.from_numpy(d["block_ids"]).pin_memory(),
                     "labels": torch.from_numpy(rng.integers(0, 40, B).astype(np.int64)).pin_memory()})
    resident = [{k: v.to(dev) for k, v in h.items()} for h in host]
    h2d = int(sum(v.numel() * v.element_size() for v in host[0].values()))
    torch.manual_seed(0)
    model = GraphConvPoolModelV1().to(dev).train()
    params = list(model.parameters())
    broadcast_parameters(params)
    opt = torch.optim.Adam(params, lr=1e-3, fused=True)
    average_grads = GradAverager(params)
    main = torch.cuda.current_stream()

    def train_step(b):
        opt.zero_grad(set_to_none=True)
        loss, _ = model(b["xyzs"], b["block_ids"], B, b["labels"])
        loss.backward()
        average_grads()
        opt.step()
        return loss

    losses = []

    def step_resident(i):
        train_step(resident[i % n_distinct])

    def step_e2e(i):
        h = host[i % n_distinct]
        loss = train_step({k: v.to(dev, non_blocking=True) for k, v in h.items()})
        buf = torch.empty((), dtype=torch.float32).pin_memory()
        buf.copy_(loss.detach(), non_blocking=True)
        main.synchronize()
        losses.append(float(buf))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = []

    def timed(fn, n, offset=0):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t0 = time.perf_counter()
        for i in range(n):
            fn(offset + i)
        e1.record()
        host_ms.append(1e3 * (time.perf_counter() - t0))
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    for i in range(max(W, n_distinct)):
        step_resident(i)
    torch.cuda.reset_peak_memory_stats()
    sampler.mark()
    profile.enabled = False
    l0 = pcs.launch_count()
    ms = timed(step_resident, K, W)
    launches = pcs.launch_count() - l0
    peak_mem = torch.cuda.max_memory_allocated()
    profile.records.clear()
    profile.enabled = True
    ms_attr = timed(step_resident, K, W)
    profile.enabled = False
    clocks = sampler.stop() if rank == 0 else None
    table = {}
    for name, a, e0, e1 in profile.records:
        tt = table.setdefault(name, {"calls": 0, "ms": 0.0, "bytes": 0, "flops": 0})
        tt["calls"] += 1
        tt["ms"] += e0.elapsed_time(e1)
        tt["bytes"] += _alg_bytes(name, a)
        tt["flops"] += _alg_flops(name, a)
    profile.records.clear()
    step_e2e(0)
    ms_e2e = timed(step_e2e, K, 1)
    assert all(np.isfinite(losses))
    pts = torch.tensor([float(B * PN * K)], device=dev)
    if world > 1:
        dist.all_reduce(pts)
        dist.barrier()
        dist.destroy_process_group()
    return dict(world=world, rank=rank, K=K, W=W, B=B, ms=ms, ms_attr=ms_attr, host_ms=host_ms[0], value=float(pts) / (ms * 1e-3),
                e2e=float(pts) / (ms_e2e * 1e-3), table=table, launches=launches, clocks=clocks, peak_mem=peak_mem, h2d=h2d,
                d2h=4, n_param=sum(p.numel() for p in params), pts_per_step=B * PN, allreduce_us=None,
                extra_config={"clouds_per_gpu_per_step": B, "points_per_cloud": PN,
                              "clouds_per_s": world * B * K / (ms * 1e-3),
                              "parallelism": f"dp{world}" + (" (one NCCL gradient all-reduce per step)" if world > 1 else ""),
                              "l2_note": "a 32-cloud step is a small working set: the step is launch-bound, not bandwidth-bound"})


def modelnet_cpu_step():
    """CPU port of the same train step (oracle layers + stock-indexing conv ops), 4 clouds per step."""
    import torch

    from oracle import cpu_layers, torch_ops
    from pointcloudsegmentation_b200.modelnet import GraphConvPoolModelV1
    from pointcloudsegmentation_b200.synthetic import modelnet_batch

    torch.manual_seed(0)
    model = GraphConvPoolModelV1(L=cpu_layers.layers(), PU=torch_ops.point_util, ops=sys.modules["oracle.torch_ops"]).train()
    d = modelnet_batch(0, 4, 1024)
    x, b, lab = torch.from_numpy(d["xyzs"]), torch.from_numpy(d["block_ids"]), torch.tensor([1, 2, 3, 4])
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)

    def step():
        opt.zero_grad(set_to_none=True)
        loss, _ = model(x, b, 4, lab)
        loss.backward()
        opt.step()
        return float(loss.detach())

    return step, int(x.shape[0])


def other_configs():
    """Short single-GPU runs of the other BASELINE.json configs, each in its own process (`bench.py --dataset X`), so
    that the driver-run line of the headline config also carries configs[1], [3] and [4]: metric, value, ms/step, e2e."""
    out = []
    for ds, steps in (("modelnet40", 10), ("scannet", 5), ("semantic3d", 5)):
        cmd = [sys.executable, str(ROOT / "bench.py"), "--dataset", ds, "--steps", str(steps), "--warmup", "3", "--no-cpu-baseline",
               "--no-ops", "--no-reference-cuda", "--no-other-configs"]
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
            d = json.loads(r.stdout.strip().splitlines()[-1])
            out.append({"dataset": ds, "metric": d["metric"], "value": d["value"], "unit": d["unit"], "ms_per_step": d["ms_per_step"],
                        "steps": d["steps"], "e2e": d["e2e"]["value"], "h2d_bytes_per_step": d["e2e"]["h2d_bytes_per_step"],
                        "d2h_bytes_per_step": d["e2e"]["d2h_bytes_per_step"], "gpu_launches": d["gpu_launches"],
                        "workload": d["config"]["workload"], "top_kernel": (d.get("roofline") or {}).get("kernel"),
                        "clocks": d.get("clocks")})
        except Exception as e:  # a failed side run must not take the headline line down with it
            out.append({"dataset": ds, "error": f"{type(e).__name__}: {e}"[:300]})
    return out


def tensor_peak():
    """Dense TF32 tensor peak = half the measured dense bf16 rate (same pipe, 2x the operand width)."""
    p = ROOT / "MEASURED_PEAKS.json"
    try:
        j = json.loads(p.read_text())
        return 0.5 * float(j.get("bf16_tflops_sustained") or j["bf16_tflops"]), "0.5 x measured dense bf16 sustained (MEASURED_PEAKS.json)"
    except Exception:
        return 0.5 * 2250.0, "0.5 x nominal dense bf16 (B200_PROFILING.md)"


def hbm_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    try:
        return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    sem = args.dataset in ("semantic3d", "modelnet40")  # the arms with their own runner / config block
    r = run_semantic3d(args) if args.dataset == "semantic3d" else (run_modelnet(args) if args.dataset == "modelnet40" else run_b200(args))
    if r["rank"] != 0:
        return 0
    peak, peak_src = hbm_peak()
    step_ms = r["ms"] / r["K"]
    rows = []
    for name, t in r["table"].items():
        gbs = (t["bytes"] / 1e9) / (t["ms"] * 1e-3) if t["ms"] > 0 else 0.0
        rows.append({"kernel": name, "calls_per_step": t["calls"] / r["K"], "ms_per_step": t["ms"] / r["K"],
                     "share_of_step": t["ms"] / r["ms_attr"], "avg_launch_us": 1e3 * t["ms"] / t["calls"],
                     "alg_bytes_per_launch": t["bytes"] / t["calls"], "achieved_gbs": gbs, "frac_of_hbm_peak": gbs / peak,
                     "achieved_fp32_tflops": (t["flops"] / 1e12) / (t["ms"] * 1e-3) if t["ms"] > 0 else 0.0})
    rows.sort(key=lambda x: -x["ms_per_step"])
    ours_ms = sum(x["ms_per_step"] for x in rows)
    # dominant kernel = the C-ABI entry with the largest share of the step
    top = rows[0] if rows else None
    traffic = traffic_src = None
    tp = ROOT / "profiles" / "top_kernel_traffic.json"  # written from an `ncu --set full` capture, per launch
    if top and tp.exists():
        try:
            tj = json.loads(tp.read_text())
            traffic = tj.get(top["kernel"])
            # NOT a measurement of this run: the capture (and the build it was taken on) is named beside the number
            traffic_src = tj.get("_source_" + top["kernel"]) or tj.get("_source")
        except Exception:
            traffic = traffic_src = None
    roofline = None
    if top:
        roofline = {"kernel": top["kernel"], "bound": "hbm", "achieved": top["achieved_gbs"], "peak": peak,
                    "unit": "GB/s", "frac": top["achieved_gbs"] / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                    "avg_launch_us": top["avg_launch_us"], "alg_bytes_per_launch": top["alg_bytes_per_launch"],
                    "share_of_step": top["share_of_step"], "all_pcs_kernels_share_of_step": ours_ms / (r["ms_attr"] / r["K"]),
                    "measured": "CUDA event pair around every launch, on the launching stream, over a second "
                                f"timed region of the same {r['K']} steps ({r['ms_attr'] / r['K']:.2f} ms/step with the events; "
                                "the headline region carries none)"}
        sm_mhz = float((r["clocks"] or {}).get("sm_mhz") or 1965.0)
        if top["kernel"] == "pcs_linear_tc":
            # tcgen05 GEMM of the point-level layers: tensor-pipe bound for the wide layers, so the roofline is
            # the tensor one; the byte figure stays beside it (the narrow layers of the same entry are HBM-bound)
            tpk, tsrc = tensor_peak()
            roofline.update({"bound": "tensor", "achieved": top["achieved_fp32_tflops"], "peak": tpk, "unit": "TFLOP/s",
                             "frac": top["achieved_fp32_tflops"] / tpk, "peak_source": tsrc,
                             "alg_flops_per_launch": top["achieved_fp32_tflops"] * 1e12 * top["avg_launch_us"] * 1e-6,
                             "mma_issue_frac": 3.0 * top["achieved_fp32_tflops"] / tpk,
                             "hbm": {"achieved": top["achieved_gbs"], "peak": peak, "unit": "GB/s", "frac": top["achieved_gbs"] / peak},
                             "note": "achieved = fp32 products of the layers (2*M*N*K) / launch time, averaged over "
                                     "the ~55 launches of a step (wide tensor-bound and narrow HBM-bound layers "
                                     "together); every product is issued as 3 TF32 MMAs (error-compensated split), so "
                                     "the tensor pipe is busy for mma_issue_frac of its TF32 peak and frac <= 1/3"})
        elif top["achieved_fp32_tflops"] > 0:
            # the fused edge-MLP kernels are fp32-FFMA bound, not byte bound: say how close to that pipe they run
            fpk = fp32_peak_tflops(sm_mhz)
            roofline["fp32_pipe"] = {"achieved": top["achieved_fp32_tflops"], "peak": fpk, "unit": "TFLOP/s",
                                     "frac": top["achieved_fp32_tflops"] / fpk,
                                     "note": "fp32-equivalent MACs of the edge MLP against the fp32 FFMA pipe it would "
                                             "otherwise run on; the kernel issues 3 TF32 MMAs per product "
                                             "(error-compensated, ~1e-6 rel.) on the tensor cores and is bound by "
                                             "HBM traffic + load latency (see the hbm fraction above)"}
        # byte movers only: the radius search is ALU-bound, the CUB sorts / scans move ~6x their algorithmic bytes
        not_movers = ("pcs_search_", "pcs_compute_permutation_info", "pcs_csr_transpose", "pcs_exclusive_scan",
                      "pcs_compute_repermutation_info", "pcs_dense_mlp3", "pcs_linear_tc", "pcs_csr_transpose_symmetric")
        movers = [x for x in rows if x["alg_bytes_per_launch"] > 0 and not x["kernel"].startswith(not_movers)]
        if movers:
            m = movers[0]
            roofline["top_hbm_bound_kernel"] = {"kernel": m["kernel"], "achieved": m["achieved_gbs"], "peak": peak,
                                                "unit": "GB/s", "frac": m["frac_of_hbm_peak"],
                                                "avg_launch_us": m["avg_launch_us"],
                                                "alg_bytes_per_launch": m["alg_bytes_per_launch"],
                                                "share_of_step": m["share_of_step"]}
    metric = {"s3dis": METRIC, "s3dis_pgnet": METRIC + " (model_pgnet tower)", "scannet": METRIC.replace("S3DIS", "ScanNet-shaped"),
              "semantic3d": "Semantic3D dense tile points/s inference",
              "modelnet40": "ModelNet40 points/s fwd+bwd"}[args.dataset]
    if sem:
        config = dict({"workload": DATASETS[args.dataset]["workload"], "points_per_gpu_per_step": r["pts_per_step"],
                       "params": r["n_param"], "host_enqueue_ms_per_step": round(r["host_ms"] / r["K"], 3),
                       "l2": f"no flush: per-step working set {r['peak_mem'] / 2**20:.0f} MiB >> 126 MB L2"}, **r["extra_config"])
        e2e = {"value": r["e2e"], "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": r["d2h"],
               "what": ("per step: the tile's xyz + features + labels from pinned host memory, block preparation + the model, "
                        "the per-centre logits copied to pinned host memory and read") if args.dataset == "semantic3d" else
                       "per step: the batch's xyz + cloud ids + labels from pinned host memory, the train step, the loss read back"}
    else:
        config = {"workload": DATASETS[args.dataset]["workload"], "blocks_per_gpu_per_step": r["B"],
                  "points_per_gpu_per_step": r["pts_per_step"], "parallelism": f"dp{r['world']} (blocks sharded)",
                  "params": r["n_param"], "host_enqueue_ms_per_step": round(r["host_ms"] / r["K"], 3),
                  "pipeline": "model.prepare (voxel pooling + 9 radius searches) of batch i+1 runs on a second "
                              "stream from a second host thread (pipeline.PrepareAhead) during step i; each "
                              "timed step = 1 prepare + 1 fwd+bwd+Adam",
                  "balance": f"every rank's batch = {r['B']} of {r['B'] + 8} candidate blocks chosen to hold the same number of "
                             "points (dist.fill_to_budget): equal work per rank per step",
                  "priming": "each of the 4 distinct synthetic batches stepped once before the warm-up steps "
                             "(cuBLAS algorithm selection and allocator growth per shape), untimed",
                  "l2": f"no flush: per-step working set {r['peak_mem'] / 2**20:.0f} MiB >> 126 MB L2"}
        e2e = {"value": r["e2e"], "unit": UNIT, "h2d_bytes_per_step": r["h2d"], "d2h_bytes_per_step": 4}
    out = {"metric": metric, "value": r["value"], "unit": UNIT, "n_gpus": r["world"], "steps": r["K"],
           "warmup": r["W"], "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config, "e2e": e2e,
           "gpu_launches": int(r["launches"]), "clocks": r["clocks"], "roofline": roofline,
           "kernels": [{k: (round(v, 4) if isinstance(v, float) else v) for k, v in x.items()} for x in rows[:12]]}
    if r["world"] == 1 and not args.no_cpu_baseline:
        use_all_host_threads()
        step, pts = (semantic3d_cpu_step() if args.dataset == "semantic3d" else modelnet_cpu_step() if args.dataset == "modelnet40"
                     else cpu_step_fn(args.cpu_blocks, dataset=args.dataset))
        step()
        n, t0 = 0, time.perf_counter()
        while n < 2 or (time.perf_counter() - t0 < 12 and n < 40):
            step(); n += 1
        dt = time.perf_counter() - t0
        sample = (f"4 clouds ({pts} points) per step, {n} steps" if args.dataset == "modelnet40" else
                  f"one 10 m x 10 m crop of a tile ({pts} dense points) per step, {n} steps" if sem else
                  f"{args.cpu_blocks} block(s)/step ({pts} points) of the same generator, {n} steps")
        out["cpu_baseline"] = {"value": pts * n / dt, "unit": UNIT, "cores": cpu_threads(), "kind": "port", "sample": sample}
    if r["allreduce_us"] is not None:
        out["allreduce_us_per_step"] = round(r["allreduce_us"], 1)  # max over ranks of the mean device time (pack + NCCL + unpack)
    if r["world"] == 1 and not args.no_reference_cuda and args.dataset in ("s3dis", "scannet"):
        ours = single_block_steps(args.dataset, "b200", 20)
        out["latency"] = {"one_block_step_ms": round(ours["ms_per_step"], 3), "points_per_step": ours["points_per_step"],
                          "value": ours["value"], "unit": UNIT,
                          "what": "fwd+bwd+Adam on ONE block per step, prepare inline on the same stream (the reference's "
                                  "unit of work: one block per tower per step, train_graph_pool.py:95-108)"}
        ref = single_block_steps(args.dataset, "reference_cuda", 5)
        if ref is None:
            out["reference_cuda"] = {"unavailable": "oracle/_ref/libref.so not built"}
        else:
            out["reference_cuda"] = dict(ref, kind="the reference's unmodified tf_ops/*.cu + PoolingIndex/*.cu kernels compiled for "
                                                   "sm_100a (oracle/_ref/libref.so) behind the same tower, literal evaluation "
                                                   "order, stock PyTorch MLPs, one block per step as the reference feeds it",
                                         b200_same_config=ours["value"], speedup_same_config=ours["value"] / ref["value"],
                                         speedup_headline=r["value"] / ref["value"])
    if r["world"] == 1 and not args.no_ops and args.dataset in ("s3dis", "scannet"):
        out["ops"] = ops_table()
    if r["world"] == 1 and args.dataset == "s3dis" and not args.no_other_configs:
        import torch

        torch.cuda.empty_cache()  # hand the cached blocks back before the side processes start
        out["other_configs"] = other_configs()
    if args.ops_report:
        Path(args.ops_report).parent.mkdir(parents=True, exist_ok=True)
        Path(args.ops_report).write_text(json.dumps({"step_ms": step_ms, "kernels": rows}, indent=1))
    print(json.dumps(out), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())

#!/usr/bin/env python
"""bench_ops.py — per-op achieved HBM GB/s vs roofline (the second half of BASELINE.json's metric),
on the op-level config: synthetic 4096-point S3DIS-shaped blocks, radius-0.1 CSR neighbours, batched
block-diagonally (default 64 blocks, so every tensor is far larger than the 126 MB L2) — plus the
same op timed in the reference's own CUDA kernel on the same GPU (oracle/_ref/libref.so) and in the
CPU oracle port on the host cores.

    python bench_ops.py [--blocks 64] [--pn 4096] [--channels 3,12,32,64,128] [--out profiles/ops.json]

Timing: CUDA events on the launching stream, 3 warm-ups, median of `--iters`; a 256 MB buffer is
written between iterations to flush L2 (and to keep the queue full, see Timer).  GB/s = algorithmic bytes / time.
"""
from __future__ import annotations

import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=64)
    ap.add_argument("--pn", type=int, default=4096)
    ap.add_argument("--channels", default="3,12,32,35,64,128,131")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--only", default="", help="comma list of op names to run")
    return ap.parse_args()


class Timer:
    """Queue [L2 flush, event, op, event] x iters WITHOUT host syncs in between, sync once at the end.
    The flush (512 MB memset, ~150 us) keeps the GPU busy while the host enqueues the next op, so the
    event pair brackets only the op's kernels — not the host's launch latency (which, timed naively,
    adds 10-30 us of GPU idle to every sub-100-us kernel)."""

    def __init__(self, iters):
        self.iters = iters
        self.flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def __call__(self, fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        evs = []
        for _ in range(self.iters):
            self.flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        return float(np.median([a.elapsed_time(b) for a, b in evs]))


def cpu_time(fn, min_s=0.5):
    fn()
    n, t0 = 0, time.perf_counter()
    while n < 2 or time.perf_counter() - t0 < min_s:
        fn(); n += 1
    return 1e3 * (time.perf_counter() - t0) / n


def main():
    args = parse()
    from oracle import libref, pcs_oracle as orc
    from pointcloudsegmentation_b200 import ops as P
    from pointcloudsegmentation_b200.synthetic import s3dis_crop

    peak = 6650.0
    try:
        peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except Exception:
        pass
    use_ref = (not args.no_ref) and libref.available()
    only = set(filter(None, args.only.split(",")))
    rng = np.random.default_rng(0)
    B, pn = args.blocks, args.pn
    pts = np.concatenate([s3dis_crop(rng, pn) for _ in range(B)])
    bb = (np.arange(B + 1) * pn).astype(np.int32)
    T = Timer(args.iters)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    dpts, dbb = dev(pts), dev(bb)
    rows = []

    def add(op, shape, alg_bytes, ms, ref_ms=None, cpu_ms=None, note=""):
        gbs = alg_bytes / 1e9 / (ms * 1e-3)
        rows.append({"op": op, "shape": shape, "alg_MB": alg_bytes / 1e6, "us": 1e3 * ms, "GBps": gbs, "frac": gbs / peak,
                     "ref_cuda_us": None if ref_ms is None else 1e3 * ref_ms,
                     "speedup_vs_ref_cuda": None if ref_ms is None else ref_ms / ms,
                     "cpu_port_ms": cpu_ms, "note": note})
        r = rows[-1]
        print(f"{op:34s} {shape:26s} {r['alg_MB']:9.1f} MB {r['us']:9.1f} us {gbs:8.1f} GB/s {100 * r['frac']:5.1f}%"
              + (f"  ref {r['ref_cuda_us']:10.1f} us x{r['speedup_vs_ref_cuda']:.1f}" if ref_ms else "")
              + (f"  cpu {cpu_ms:9.1f} ms" if cpu_ms else ""), flush=True)

    want = lambda n: (not only) or n in only  # noqa: E731

    # ---- radius search (a18) and the CSR it yields -------------------------------------------------
    r2 = 0.1 * 0.1
    nidxs, nlens, nbegs, ncens = P.search_neighborhood_brute_force(dpts, r2, block_begs=dbb, max_block_pn=pn)
    P_all, en = B * pn, int(nidxs.shape[0])
    print(f"# {B} blocks x {pn} pts = {P_all} points, {en} edges (mean degree {en / P_all:.2f}); HBM peak {peak} GB/s")
    if want("search"):
        ms = T(lambda: P.search_neighborhood_brute_force(dpts, r2, block_begs=dbb, max_block_pn=pn))
        ref_ms = None
        if use_ref:  # the reference searches one block per call
            one = dpts[:pn].contiguous()
            ref_ms = T(lambda: libref.search(one, float(np.float32(r2)))) * B
        cpu_ms = None if args.no_cpu else cpu_time(lambda: orc.search(pts[:pn], r2)) * B
        add("search_neighborhood (count+fill)", f"{B}x{pn} r=0.1", 12 * P_all + 8 * en + 8 * P_all, ms, ref_ms, cpu_ms,
            f"{2 * B * pn * pn / (ms * 1e-3) / 1e12:.2f} T candidate tests/s")
        # the same search on voxel-sorted points (what points_pooling hands to the networks): the chunk
        # culling only helps when consecutive indices are spatially coherent
        bids = torch.arange(B, device="cuda", dtype=torch.int32).repeat_interleave(pn)
        o2p = P.compute_permutation_info(P.compute_voxel_index(dpts, 0.15, 3.0), block_ids=bids)[0]
        spts = P.permutate_feature(dpts, o2p)
        ms2 = T(lambda: P.search_neighborhood_brute_force(spts, r2, block_begs=dbb, max_block_pn=pn))
        add("search_neighborhood, voxel-sorted", f"{B}x{pn} r=0.1", 12 * P_all + 8 * en + 8 * P_all, ms2, ref_ms, cpu_ms,
            f"{2 * B * pn * pn / (ms2 * 1e-3) / 1e12:.2f} T brute-force-equivalent tests/s")
    hn, hl, hb, hc = (t.cpu().numpy() for t in (nidxs, nlens, nbegs, ncens))
    P.csr_transpose(nidxs, P_all)

    for C in [int(c) for c in args.channels.split(",")]:
        f_pt = torch.rand(P_all, C, device="cuda") * 2 - 1
        f_ed = torch.rand(en, C, device="cuda") * 2 - 1
        hf_pt = hf_ed = None
        if not args.no_cpu:
            hf_pt, hf_ed = f_pt.cpu().numpy(), f_ed.cpu().numpy()
        b_sc = 4 * C * (P_all + en) + 4 * en + 8 * P_all
        b_sg = 4 * C * (P_all + en) + 8 * P_all
        b_ss = 4 * C * (P_all + en) + 4 * en
        b_mg = 4 * C * en + 8 * C * P_all + 8 * P_all
        shape = f"P={P_all} E={en} C={C}"
        for ud in (False, True):
            if want("neighbor_scatter"):
                add(f"neighbor_scatter(diff={int(ud)})", shape, b_sc,
                    T(lambda: P.neighbor_scatter(f_pt, nidxs, nlens, nbegs, ud)),
                    T(lambda: libref.neighbor_scatter(f_pt, nidxs, nlens, nbegs, ud)) if use_ref else None,
                    None if args.no_cpu else cpu_time(lambda: orc.neighbor_scatter(hf_pt, hn, hl, hb, ud)))
            if want("neighbor_gather"):
                add(f"neighbor_gather(diff={int(ud)})", shape, b_sc,
                    T(lambda: P.neighbor_gather(f_ed, nidxs, nlens, nbegs, ud)),
                    T(lambda: libref.neighbor_gather(f_ed, nidxs, nlens, nbegs, ud)) if use_ref else None,
                    None if args.no_cpu else cpu_time(lambda: orc.neighbor_gather(hf_ed, hn, hl, hb, ud)))
        if want("sum_feat_gather"):
            add("neighbor_sum_feat_gather", shape, b_sg, T(lambda: P.neighbor_sum_feat_gather(f_ed, ncens, nlens, nbegs)),
                T(lambda: libref.sum_feat_gather(f_ed, nlens, nbegs)) if use_ref else None,
                None if args.no_cpu else cpu_time(lambda: orc.segment_sum(hf_ed, hl, hb)))
        if want("sum_feat_scatter"):
            add("neighbor_sum_feat_scatter", shape, b_ss, T(lambda: P.neighbor_sum_feat_scatter(f_pt, ncens, nlens, nbegs)),
                T(lambda: libref.sum_feat_scatter(f_pt, ncens)) if use_ref else None,
                None if args.no_cpu else cpu_time(lambda: orc.row_gather(hf_pt, hc)))
        if want("max_feat_gather"):
            g, mi = P.neighbor_max_feat_gather(f_ed, nlens, nbegs)
            add("neighbor_max_feat_gather", shape, b_mg, T(lambda: P.neighbor_max_feat_gather(f_ed, nlens, nbegs)),
                T(lambda: libref.max_feat_gather(f_ed, nlens, nbegs)) if use_ref else None,
                None if args.no_cpu else cpu_time(lambda: orc.max_feat_gather(hf_ed, hl, hb)))
            hmi = mi.cpu().numpy()
            add("neighbor_max_feat_scatter", shape, b_mg, T(lambda: P.neighbor_max_feat_scatter(g, en, mi, nbegs, nlens)),
                T(lambda: libref.max_feat_scatter(g, mi, nbegs, nlens, en)) if use_ref else None,
                None if args.no_cpu else cpu_time(lambda: orc.max_feat_scatter(hf_pt, hmi, hb, hl, en)))
        if want("permutate_feature"):
            perm = torch.randperm(P_all, device="cuda").to(torch.int32)
            hperm = perm.cpu().numpy()
            add("permutate_feature", f"P={P_all} C={C}", 8 * C * P_all + 4 * P_all, T(lambda: P.permutate_feature(f_pt, perm)),
                T(lambda: libref.permutate_feature(f_pt, perm)) if use_ref else None,
                None if args.no_cpu else cpu_time(lambda: orc.row_gather(hf_pt, hperm)))
        del f_pt, f_ed

    # ---- location-weight ops (a7-a10), m = 26 anchors -------------------------------------------
    m = 26
    if want("location_weight"):
        lw = torch.rand(en, m, device="cuda")
        dl = torch.rand(P_all, m, device="cuda")
        hlw = None if args.no_cpu else lw.cpu().numpy()
        b = 4 * m * (en + P_all) + 8 * P_all
        add("location_weight_sum", f"E={en} m={m}", b, T(lambda: P.location_weight_sum(lw, nlens, nbegs)),
            T(lambda: libref.locw_sum(lw, nlens, nbegs)) if use_ref else None,
            None if args.no_cpu else cpu_time(lambda: orc.segment_sum(hlw, hl, hb)))
        add("location_weight_sum_backward", f"E={en} m={m}", b, T(lambda: P.location_weight_sum_backward(lw, dl, nlens, nbegs)),
            T(lambda: libref.locw_sum_backward(dl, nlens, nbegs, en)) if use_ref else None, None)
        for ofn in (8, 32):
            Bq = max(1, B // (4 if ofn == 32 else 1))  # keep the [E,m,ofn] tensor within a few GB
            e2, p2 = int(nbegs[Bq * pn - 1] + nlens[Bq * pn - 1]), Bq * pn
            l2, b2 = nlens[:p2].contiguous(), nbegs[:p2].contiguous()
            lw2 = lw[:e2].contiguous()
            tf = torch.rand(e2, m, ofn, device="cuda") * 2 - 1
            d = torch.rand(p2, m, ofn, device="cuda") * 2 - 1
            bf = 4 * e2 * m * (ofn + 1) + 4 * p2 * m * ofn + 8 * p2
            bb_ = 8 * e2 * m * (ofn + 1) + 4 * p2 * m * ofn + 8 * p2
            add("location_weight_feat_sum", f"E={e2} m={m} ofn={ofn}", bf, T(lambda: P.location_weight_feat_sum(lw2, tf, l2, b2)),
                T(lambda: libref.locw_feat_sum(lw2, tf, l2, b2)) if use_ref else None, None)
            add("location_weight_feat_sum_backward", f"E={e2} m={m} ofn={ofn}", bb_,
                T(lambda: P.location_weight_feat_sum_backward(lw2, tf, d, l2, b2)),
                T(lambda: libref.locw_feat_sum_backward(lw2, tf, d, l2, b2)) if use_ref else None, None)
            del tf, d

    # ---- pooling index pipeline (a12-a17) and eliminate_center (a11) -----------------------------
    if want("pooling"):
        bids = torch.arange(B, device="cuda", dtype=torch.int32).repeat_interleave(pn)
        vidx = P.compute_voxel_index(dpts, 0.15, 3.0)
        add("compute_voxel_index", f"P={P_all}", 24 * P_all, T(lambda: P.compute_voxel_index(dpts, 0.15, 3.0)),
            None, None if args.no_cpu else cpu_time(lambda: orc.voxel_index(pts, 0.15, 3.0)))
        hv, hbids = vidx.cpu().numpy(), bids.cpu().numpy()
        ref_ms = None
        if use_ref:
            v1 = vidx[:pn].contiguous()
            ref_ms = T(lambda: libref.permutation_info(v1)) * B
        add("compute_permutation_info", f"P={P_all} ({B} blocks)", 32 * P_all, T(lambda: P.compute_permutation_info(vidx, block_ids=bids)),
            ref_ms, None if args.no_cpu else cpu_time(lambda: orc.permutation_info(hv, None, hbids)), "includes the D2H size sync")
        add("eliminate_center", f"P={P_all} E={en}", 12 * en + 16 * P_all - 8 * P_all, T(lambda: P.eliminate_center(nidxs, nlens, nbegs)),
            T(lambda: libref.eliminate_center(nidxs, nlens, nbegs)) if use_ref else None,
            None if args.no_cpu else cpu_time(lambda: orc.eliminate_center(hn, hl, hb)))

    out = {"gpu": torch.cuda.get_device_name(0), "hbm_peak_gbs": peak, "blocks": B, "pn": pn, "points": P_all, "edges": en,
           "cpu_threads": orc.num_threads(), "rows": rows}
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(out, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())

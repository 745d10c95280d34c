#!/usr/bin/env python
"""bench_ops.py — per-op achieved HBM GB/s vs roofline (the second half of BASELINE.json's metric) on the
op-level config (`configs[0]`): synthetic 4096-point S3DIS-shaped blocks, radius-0.1 CSR neighbours
(mean degree 12), batched block-diagonally — with the same op timed in the reference's own CUDA kernel on
the same GPU (oracle/_ref/libref.so, "the kernel to beat") and, optionally, in the CPU oracle port.

    python bench_ops.py [--blocks 64] [--pn 4096] [--channels 3,12,32,35,64,128,131] [--dtypes f32,bf16]
                        [--out profiles/ops.json]

Sizes.  A 64-block batch at C = 3 is a 55 MB tensor: 8 µs at the HBM rate, so launch ramp and tail (2-4 µs)
would cap ANY kernel near 50-60 %.  The roofline figure is therefore taken on a batch whose edge tensor
is >= `--min-mb` (default 768 MB of algorithmic bytes, >> the 126 MB L2) — `--blocks` is the floor, narrow
widths get proportionally more 4096-point blocks of the same generator (SURVEY §8d "batched variant").  The
block count used is printed in every row.  Points are voxel-sorted per block, as `points_pooling` hands
them to every network (the reference stores its hierarchies sorted the same way).

Timing: CUDA events on the launching stream, 3 warm-ups, median of `--iters`; a 512 MB buffer is written
between iterations (L2 flush; it also keeps the queue full, see Timer).  GB/s = algorithmic bytes
(every input element read once, every output element written once) / time.

`run_table()` is what bench.py calls for the `ops` key of the driver-run line.
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

DEFAULT_CHANNELS = (3, 12, 32, 35, 64, 128, 131)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=64)
    ap.add_argument("--pn", type=int, default=4096)
    ap.add_argument("--channels", default=",".join(map(str, DEFAULT_CHANNELS)))
    ap.add_argument("--dtypes", default="f32,bf16")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--min-mb", type=float, default=768.0)
    ap.add_argument("--max-blocks", type=int, default=1024)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--cpu", action="store_true", help="also time the CPU oracle port (slow)")
    ap.add_argument("--only", default="", help="comma list of op groups to run")
    ap.add_argument("--quiet", action="store_true")
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


def hbm_peak():
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def run_table(blocks=64, pn=4096, channels=DEFAULT_CHANNELS, dtypes=("f32", "bf16"), iters=10, min_mb=768.0,
              max_blocks=1024, use_ref=True, use_cpu=False, only=(), log=print):
    """-> dict(rows=[...], ...).  Every row: op, dtype, C, blocks, alg_MB, us, GBps, frac, ref_cuda_us."""
    from oracle import libref, pcs_oracle as orc
    from pointcloudsegmentation_b200 import ops as P
    from pointcloudsegmentation_b200.synthetic import s3dis_crop

    peak, peak_src = hbm_peak()
    use_ref = use_ref and libref.available()
    only = set(only)
    want = lambda n: (not only) or n in only  # noqa: E731
    rng = np.random.default_rng(0)
    dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
    T = Timer(iters)
    rows = []
    tdt = {"f32": torch.float32, "bf16": torch.bfloat16}
    esz = {"f32": 4, "bf16": 2}
    bf16_ok = hasattr(P, "FEATURE_DTYPES") and torch.bfloat16 in P.FEATURE_DTYPES
    dtypes = [d for d in dtypes if d == "f32" or bf16_ok]

    def add(op, dt, C, nb, shape, alg_bytes, ms, ref_ms=None, cpu_ms=None, note=""):
        gbs = alg_bytes / 1e9 / (ms * 1e-3)
        r = {"op": op, "dtype": dt, "C": C, "blocks": nb, "shape": shape, "alg_MB": alg_bytes / 1e6, "us": 1e3 * ms,
             "GBps": gbs, "frac": gbs / peak, "ref_cuda_us": None if ref_ms is None else 1e3 * ref_ms,
             "speedup_vs_ref_cuda": None if ref_ms is None else ref_ms / ms, "cpu_port_ms": cpu_ms, "note": note}
        rows.append(r)
        log(f"{op:34s} {dt:4s} {shape:30s} {r['alg_MB']:9.1f} MB {r['us']:9.1f} us {gbs:8.1f} GB/s {100 * r['frac']:5.1f}%"
            + (f"  ref {r['ref_cuda_us']:10.1f} us x{r['speedup_vs_ref_cuda']:.1f}" if ref_ms else "")
            + (f"  cpu {cpu_ms:9.1f} ms" if cpu_ms else ""))

    # ---- the batch: B_max blocks, voxel-sorted per block (what points_pooling hands to every layer) ----
    per_block_mb = lambda C, e: e * C * pn * 13 / 1e6  # noqa: E731  (edge + point tensor of one block, MB; degree 12)
    need = lambda C, e: int(min(max_blocks, max(blocks, np.ceil(min_mb / per_block_mb(C, e)))))  # noqa: E731
    B_max = max([need(C, esz[d]) for C in channels for d in dtypes] + [blocks])
    pts = np.concatenate([s3dis_crop(rng, pn) for _ in range(min(B_max, 64))])
    if B_max > 64:  # more blocks = the same 64 geometries again (block-diagonal: they never interact)
        pts = np.tile(pts, ((B_max + 63) // 64, 1))[:B_max * pn]
    bb = (np.arange(B_max + 1) * pn).astype(np.int32)
    dbb = dev(bb)
    bids = torch.arange(B_max, device="cuda", dtype=torch.int32).repeat_interleave(pn)
    raw = dev(pts)
    o2p = P.compute_permutation_info(P.compute_voxel_index(raw, 0.15, 3.0), block_ids=bids)[0]
    dpts = P.permutate_feature(raw, o2p)
    r2 = 0.1 * 0.1
    nidxs, nlens, nbegs, ncens = P.search_neighborhood_brute_force(dpts, r2, block_begs=dbb, max_block_pn=pn)
    en_all = int(nidxs.shape[0])
    log(f"# {B_max} blocks x {pn} pts = {B_max * pn} points, {en_all} edges (mean degree {en_all / (B_max * pn):.2f}); "
        f"HBM peak {peak} GB/s ({peak_src}); rows use the first `blocks` blocks of this batch")

    def prefix(nb):
        """CSR of the first nb blocks (a prefix of a block-diagonal CSR is a CSR)."""
        p = nb * pn
        e = int(nbegs[p - 1] + nlens[p - 1])
        return p, e, nidxs[:e].contiguous(), nlens[:p].contiguous(), nbegs[:p].contiguous(), ncens[:e].contiguous()

    # ---- radius search (a18) ------------------------------------------------------------------------
    if want("search"):
        B = blocks
        p, e = B * pn, int(nbegs[B * pn - 1] + nlens[B * pn - 1])
        sp, sb = dpts[:p].contiguous(), dbb[:B + 1].contiguous()
        ms2 = T(lambda: P.search_neighborhood_brute_force(sp, r2, block_begs=sb, max_block_pn=pn))
        ref_ms = cpu_ms = None
        if use_ref:  # the reference searches one block per call
            one = dpts[:pn].contiguous()
            ref_ms = T(lambda: libref.search(one, float(np.float32(r2)))) * B
        if use_cpu:
            h = dpts[:pn].cpu().numpy()
            cpu_ms = cpu_time(lambda: orc.search(h, r2)) * B
        add("search_neighborhood (voxel-sorted)", "f32", 3, B, f"{B}x{pn} r=0.1", 20 * p + 8 * e, ms2, ref_ms, cpu_ms,
            f"ALU-bound: {2 * B * pn * pn / (ms2 * 1e-3) / 1e12:.2f} T brute-force-equivalent tests/s")
        rp = raw[:p].contiguous()
        ms = T(lambda: P.search_neighborhood_brute_force(rp, r2, block_begs=sb, max_block_pn=pn))
        add("search_neighborhood (random order)", "f32", 3, B, f"{B}x{pn} r=0.1", 20 * p + 8 * e, ms, ref_ms, cpu_ms)

    for dt in dtypes:
        s = esz[dt]
        for C in channels:
            nb = need(C, s)
            P_all, en, ni, nl, nbg, nc = prefix(nb)
            P.csr_transpose(ni, P_all)
            f_pt = (torch.rand(P_all, C, device="cuda") * 2 - 1).to(tdt[dt])
            f_ed = (torch.rand(en, C, device="cuda") * 2 - 1).to(tdt[dt])
            ref = use_ref and dt == "f32"
            cpu = use_cpu and dt == "f32" and nb <= 64
            if cpu:
                hf_pt, hf_ed = f_pt.cpu().numpy(), f_ed.cpu().numpy()
                hn, hl, hb, hc = (t.cpu().numpy() for t in (ni, nl, nbg, nc))
            b_sc = s * C * (P_all + en) + 4 * en + 8 * P_all
            b_sg = s * C * (P_all + en) + 8 * P_all
            b_ss = s * C * (P_all + en) + 4 * en
            b_mg = s * C * en + (s + 4) * C * P_all + 8 * P_all
            shape = f"P={P_all} E={en} C={C}"
            for ud in (False, True):
                if want("neighbor_scatter"):
                    add(f"neighbor_scatter(diff={int(ud)})", dt, C, nb, shape, b_sc,
                        T(lambda: P.neighbor_scatter(f_pt, ni, nl, nbg, ud)),
                        T(lambda: libref.neighbor_scatter(f_pt, ni, nl, nbg, ud)) if ref else None,
                        cpu_time(lambda: orc.neighbor_scatter(hf_pt, hn, hl, hb, ud)) if cpu else None)
                if want("neighbor_gather"):
                    add(f"neighbor_gather(diff={int(ud)})", dt, C, nb, shape, b_sc,
                        T(lambda: P.neighbor_gather(f_ed, ni, nl, nbg, ud)),
                        T(lambda: libref.neighbor_gather(f_ed, ni, nl, nbg, ud)) if ref else None,
                        cpu_time(lambda: orc.neighbor_gather(hf_ed, hn, hl, hb, ud)) if cpu else None)
            if want("sum_feat_gather"):
                add("neighbor_sum_feat_gather", dt, C, nb, shape, b_sg, T(lambda: P.neighbor_sum_feat_gather(f_ed, nc, nl, nbg)),
                    T(lambda: libref.sum_feat_gather(f_ed, nl, nbg)) if ref else None,
                    cpu_time(lambda: orc.segment_sum(hf_ed, hl, hb)) if cpu else None)
            if want("sum_feat_scatter"):
                add("neighbor_sum_feat_scatter", dt, C, nb, shape, b_ss, T(lambda: P.neighbor_sum_feat_scatter(f_pt, nc, nl, nbg)),
                    T(lambda: libref.sum_feat_scatter(f_pt, nc)) if ref else None,
                    cpu_time(lambda: orc.row_gather(hf_pt, hc)) if cpu else None)
            if want("max_feat_gather"):
                g, mi = P.neighbor_max_feat_gather(f_ed, nl, nbg)
                add("neighbor_max_feat_gather", dt, C, nb, shape, b_mg, T(lambda: P.neighbor_max_feat_gather(f_ed, nl, nbg)),
                    T(lambda: libref.max_feat_gather(f_ed, nl, nbg)) if ref else None,
                    cpu_time(lambda: orc.max_feat_gather(hf_ed, hl, hb)) if cpu else None)
                add("neighbor_max_feat_scatter", dt, C, nb, shape, b_mg, T(lambda: P.neighbor_max_feat_scatter(g, en, mi, nbg, nl)),
                    T(lambda: libref.max_feat_scatter(g, mi, nbg, nl, en)) if ref else None, None)
                del g, mi
            if want("permutate_feature"):
                # the permutation points_pooling applies: a voxel sort of randomly ordered points.  A point tensor is
                # 13x smaller than the edge tensor of the same batch, so the batch is repeated (same permutation,
                # shifted per repeat) until this row, too, moves >= min_mb
                perm = o2p[:P_all].contiguous()
                rep = int(max(1, np.ceil(min_mb * 1e6 / (2 * s * C * P_all + 4 * P_all))))
                if rep > 1:
                    perm = (perm.unsqueeze(0) + (torch.arange(rep, device="cuda", dtype=torch.int32) * P_all).unsqueeze(1)).reshape(-1)
                    f_rep = f_pt.repeat(rep, 1)
                else:
                    f_rep = f_pt
                Pr = P_all * rep
                add("permutate_feature", dt, C, nb * rep, f"P={Pr} C={C}", 2 * s * C * Pr + 4 * Pr,
                    T(lambda: P.permutate_feature(f_rep, perm)),
                    T(lambda: libref.permutate_feature(f_rep, perm)) if ref else None,
                    cpu_time(lambda: orc.row_gather(hf_pt, perm.cpu().numpy())) if cpu and rep == 1 else None)
                del f_rep, perm
            del f_pt, f_ed

        # ---- location-weight ops (a7-a10), m = 26 anchors ---------------------------------------------
        m = 26
        if want("location_weight"):
            nb = need(m, s)
            P_all, en, ni, nl, nbg, nc = prefix(nb)
            ref = use_ref and dt == "f32"
            lw = torch.rand(en, m, device="cuda").to(tdt[dt])
            dl = torch.rand(P_all, m, device="cuda").to(tdt[dt])
            b = s * m * (en + P_all) + 8 * P_all
            add("location_weight_sum", dt, m, nb, f"E={en} m={m}", b, T(lambda: P.location_weight_sum(lw, nl, nbg)),
                T(lambda: libref.locw_sum(lw, nl, nbg)) if ref else None, None)
            add("location_weight_sum_backward", dt, m, nb, f"E={en} m={m}", b, T(lambda: P.location_weight_sum_backward(lw, dl, nl, nbg)),
                T(lambda: libref.locw_sum_backward(dl, nl, nbg, en)) if ref else None, None)
            del lw, dl
            for ofn in (8, 32):
                Bq = max(1, blocks // (4 if ofn == 32 else 1))  # keep the [E,m,ofn] tensor within a few GB
                p2, e2, _, l2, b2, _ = prefix(Bq)
                lw2 = torch.rand(e2, m, device="cuda").to(tdt[dt])
                tf = (torch.rand(e2, m, ofn, device="cuda") * 2 - 1).to(tdt[dt])
                d = (torch.rand(p2, m, ofn, device="cuda") * 2 - 1).to(tdt[dt])
                bf = s * e2 * m * (ofn + 1) + s * p2 * m * ofn + 8 * p2
                bb_ = 2 * s * e2 * m * (ofn + 1) + s * p2 * m * ofn + 8 * p2
                add("location_weight_feat_sum", dt, m * ofn, Bq, f"E={e2} m={m} ofn={ofn}", bf, T(lambda: P.location_weight_feat_sum(lw2, tf, l2, b2)),
                    T(lambda: libref.locw_feat_sum(lw2, tf, l2, b2)) if ref else None, None)
                add("location_weight_feat_sum_backward", dt, m * ofn, Bq, f"E={e2} m={m} ofn={ofn}", bb_,
                    T(lambda: P.location_weight_feat_sum_backward(lw2, tf, d, l2, b2)),
                    T(lambda: libref.locw_feat_sum_backward(lw2, tf, d, l2, b2)) if ref else None, None)
                del tf, d, lw2

    # ---- pooling index pipeline (a12-a17) and eliminate_center (a11): integer / coordinate streams ----
    if want("pooling"):
        nb = B_max
        P_all, en, ni, nl, nbg, nc = prefix(nb)
        rp = raw[:P_all].contiguous()
        vidx = P.compute_voxel_index(rp, 0.15, 3.0)
        bid = bids[:P_all].contiguous()
        ref_ms = None
        if use_ref:
            v1 = vidx[:pn].contiguous()
            ref_ms = T(lambda: libref.permutation_info(v1)) * nb
        add("compute_permutation_info", "i32", 3, nb, f"P={P_all} ({nb} blocks)", 32 * P_all,
            T(lambda: P.compute_permutation_info(vidx, block_ids=bid)), ref_ms, None,
            "CUB radix sort: moves ~6x the algorithmic bytes; includes the D2H size sync")
        add("eliminate_center", "i32", 1, nb, f"P={P_all} E={en}", 12 * en + 8 * P_all, T(lambda: P.eliminate_center(ni, nl, nbg)),
            T(lambda: libref.eliminate_center(ni, nl, nbg)) if use_ref else None, None)
        # coordinate streams (24-28 bytes per point): the batch repeated until the row moves >= min_mb
        rep = int(max(1, np.ceil(min_mb * 1e6 / (24 * P_all))))
        rp = rp.repeat(rep, 1)
        Pr = P_all * rep
        add("compute_voxel_index", "f32", 3, nb * rep, f"P={Pr}", 24 * Pr, T(lambda: P.compute_voxel_index(rp, 0.15, 3.0)),
            T(lambda: libref.voxel_index(rp, 0.15, -1.501, -1.501, -0.001)) if use_ref else None, None)
        cx = torch.rand(Pr // 8 + 1, 3, device="cuda")
        ci = torch.randint(0, Pr // 8 + 1, (Pr,), device="cuda", dtype=torch.int32).sort().values.contiguous()
        add("compute_diff_xyz", "f32", 3, nb * rep, f"P={Pr}", 28 * Pr + 12 * (Pr // 8 + 1), T(lambda: P.compute_diff_xyz(rp, cx, ci)),
            T(lambda: libref.diff_xyz(rp, cx, ci)) if use_ref else None, None)

    feature_rows = [r for r in rows if not r["op"].startswith(("search", "compute_permutation_info"))]
    below = sorted({f"{r['op']}[{r['dtype']},C={r['C']}]" for r in feature_rows if r["frac"] < 0.70})
    return {"gpu": torch.cuda.get_device_name(0), "hbm_peak_gbs": peak, "hbm_peak_source": peak_src, "floor_blocks": blocks,
            "pn": pn, "min_alg_mb": min_mb, "iters": iters, "cpu_threads": orc.num_threads(),
            "min_frac_over_ops": min((r["frac"] for r in feature_rows), default=None),
            "ops_below_70pct": below, "rows": rows}


def main():
    args = parse()
    out = run_table(args.blocks, args.pn, [int(c) for c in args.channels.split(",")], args.dtypes.split(","), args.iters,
                    args.min_mb, args.max_blocks, not args.no_ref, args.cpu, set(filter(None, args.only.split(","))),
                    (lambda *a: None) if args.quiet else (lambda s: print(s, flush=True)))
    print(f"# min frac over feature ops {out['min_frac_over_ops']:.3f}; below 70 %: {len(out['ops_below_70pct'])} "
          f"of {len(out['rows'])} rows", flush=True)
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text(json.dumps(out, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())

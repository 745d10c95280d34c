#!/usr/bin/env python
"""Generate tests/golden/*.npz by EXECUTING the reference's own numpy test oracles.

Run in the authoring container only (needs /root/reference; the GPU box has none):

    python tests/golden/make_golden.py

The reference's tests are Python-2 scripts that import TensorFlow at module level, so they cannot be
imported.  Instead the named function definitions are located with `ast` in the reference files where
they lie, compiled from that source text and executed here (nothing is copied into the repo), with
the two py2-isms they rely on supplied: `xrange` -> range, `np.int` -> int; for
np_concat_non_center_backward the py2 integer `/` is mapped to `//` (ast.Div -> ast.FloorDiv).
Inputs follow the reference tests' own generators (random CSR with empty rows, U(-1,1) features:
tf_ops/test/test_neighbor_scatter.py:171-185, test_feats_sum_scatter.py:52-60, test_max_pool.py:64-72)
but seeded.  Outputs are float64 exactly as the reference functions return them.
"""
from __future__ import annotations

import ast
import sys
from pathlib import Path

import numpy as np

REF = Path("/root/reference/tf_ops/test")
OUT = Path(__file__).resolve().parent

WANTED = {
    "test_neighbor_scatter.py": ["copy_scatter_np", "copy_gather_np", "diff_scatter_np", "diff_gather_np"],
    "test_feats_sum_scatter.py": ["sum_scatter_forward_np", "sum_scatter_backward_np"],
    "np_ops.py": ["np_pool_forward", "np_pool_backward", "np_unpool_forward", "np_unpool_backward",
                  "np_concat_non_center_forward", "np_concat_non_center_backward"],
    "test_permutate.py": ["permutate_forward"],
    "data_util.py": ["compute_nidxs_bgs", "compute_cidxs"],
}


class _Py2Div(ast.NodeTransformer):
    def visit_BinOp(self, node):
        self.generic_visit(node)
        if isinstance(node.op, ast.Div):
            node.op = ast.FloorDiv()
        return node


def load_reference_functions() -> dict:
    if not hasattr(np, "int"):
        np.int = int  # removed in numpy >= 1.24; the reference uses np.int as a dtype
    ns: dict = {"np": np, "xrange": range}
    for fname, names in WANTED.items():
        src = (REF / fname).read_text()
        # the files are py2 (print statements): parse function by function from their source segments
        lines = src.splitlines(keepends=True)
        starts = [i for i, l in enumerate(lines) if l.startswith("def ")]
        for name in names:
            s = next(i for i in starts if lines[i].startswith(f"def {name}("))
            e = s + 1
            while e < len(lines) and (lines[e].startswith((" ", "\t")) or lines[e].strip() == ""):
                e += 1
            tree = ast.parse("".join(lines[s:e]))
            if name == "np_concat_non_center_backward":
                tree = ast.fix_missing_locations(_Py2Div().visit(tree))
            exec(compile(tree, str(REF / fname), "exec"), ns)
    return ns


def random_csr(rng, pn, max_len, min_len=0):
    """lens in [min_len,max_len), neighbours without replacement (test_neighbor_scatter.py:171-185)."""
    lens = rng.integers(min_len, max_len, pn)
    nidxs = [rng.choice(pn, int(n), replace=False) for n in lens if n > 0]
    nidxs = np.concatenate(nidxs) if nidxs else np.zeros(0, np.int64)
    return nidxs.astype(np.int32), lens.astype(np.int32)


def main() -> int:
    if not REF.exists():
        print("no /root/reference here: golden fixtures are generated in the authoring container", file=sys.stderr)
        return 1
    R = load_reference_functions()
    rng = np.random.default_rng(20181)

    # ---- a1/a2 neighbor_scatter / neighbor_gather, copy and diff --------------------------------
    for tag, (pn, fd) in {"small": (57, 33), "wide": (40, 130), "odd": (101, 7)}.items():
        nidxs, lens = random_csr(rng, pn, 10)
        bgs = R["compute_nidxs_bgs"](lens)
        cens = R["compute_cidxs"](lens)
        feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
        dsfeats = rng.uniform(-1, 1, (nidxs.shape[0], fd)).astype(np.float32)
        f64, d64 = feats.astype(np.float64), dsfeats.astype(np.float64)
        np.savez_compressed(
            OUT / f"neighbor_scatter_{tag}.npz", feats=feats, dsfeats=dsfeats, nidxs=nidxs, lens=lens,
            bgs=bgs.astype(np.int32), cens=cens,
            copy_fwd=R["copy_scatter_np"](f64, nidxs, lens, bgs, cens),
            copy_bwd=R["copy_gather_np"](f64, d64, nidxs, lens, bgs, cens),
            diff_fwd=R["diff_scatter_np"](f64, nidxs, lens, bgs, cens),
            diff_bwd=R["diff_gather_np"](f64, d64, nidxs, lens, bgs, cens))

    # ---- a4/a3 sum scatter fwd / bwd (= sum gather) ------------------------------------------------
    for tag, (pn, fd) in {"small": (64, 20), "wide": (31, 257)}.items():
        lens = rng.integers(0, 5, pn).astype(np.int32)
        bgs = R["compute_nidxs_bgs"](lens)
        cens = R["compute_cidxs"](lens)
        feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
        dsfeats = rng.uniform(-1, 1, (cens.shape[0], fd)).astype(np.float32)
        np.savez_compressed(
            OUT / f"sum_scatter_{tag}.npz", feats=feats, dsfeats=dsfeats, lens=lens, bgs=bgs.astype(np.int32),
            cens=cens, fwd=R["sum_scatter_forward_np"](feats.astype(np.float64), lens, bgs, cens),
            bwd=R["sum_scatter_backward_np"](feats.astype(np.float64), dsfeats.astype(np.float64), lens, bgs, cens))

    # ---- a5/a6 max pool fwd (values + argmax) / bwd; unpool fwd/bwd ----------------------------------
    for tag, (pn2, fd, lo, hi) in {"pool": (30, 19, 3, 5), "ragged": (45, 64, 1, 10)}.items():
        vlens = rng.integers(lo, hi, pn2).astype(np.int32)
        vbegs = R["compute_nidxs_bgs"](vlens)
        vcens = R["compute_cidxs"](vlens)
        pn1 = int(vlens.sum())
        ifeats = rng.uniform(-1, 1, (pn1, fd)).astype(np.float32)
        i64 = ifeats.astype(np.float64)
        pfeats, max_idxs = R["np_pool_forward"](i64, vlens, vbegs)
        dpfeats = rng.uniform(-1, 1, pfeats.shape).astype(np.float32)
        difeats = R["np_pool_backward"](dpfeats.astype(np.float64), pfeats, i64, vlens, vbegs)
        up_in = rng.uniform(-1, 1, (pn2, fd)).astype(np.float32)
        up = R["np_unpool_forward"](up_in.astype(np.float64), vlens, vbegs, vcens)
        dup = rng.uniform(-1, 1, up.shape).astype(np.float32)
        dup_in = R["np_unpool_backward"](dup.astype(np.float64), up, up_in.astype(np.float64), vlens, vbegs, vcens)
        np.savez_compressed(
            OUT / f"max_pool_{tag}.npz", ifeats=ifeats, vlens=vlens, vbegs=vbegs.astype(np.int32), vcens=vcens,
            pfeats=pfeats, max_idxs=max_idxs.astype(np.int32), dpfeats=dpfeats, difeats=difeats,
            up_in=up_in, up=up, dup=dup, dup_in=dup_in)

    # ---- a11 composite: concat non-centre scatter fwd/bwd (np_ops.py:79-114) -------------------------
    pn, fd = 40, 9
    pts = rng.uniform(0, 1, (pn, 3))
    d2 = ((pts[:, None, :] - pts[None, :, :]) ** 2).sum(-1)
    rows = [np.nonzero(d2[i] < 0.09)[0] for i in range(pn)]  # contains i itself
    lens = np.array([len(r) for r in rows], np.int32)
    nidxs = np.concatenate(rows).astype(np.int32)
    bgs = R["compute_nidxs_bgs"](lens)
    feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
    fwd = R["np_concat_non_center_forward"](feats.astype(np.float64), nidxs, lens, bgs)
    dfwd = rng.uniform(-1, 1, fwd.shape).astype(np.float32)
    bwd = R["np_concat_non_center_backward"](dfwd.astype(np.float64), nidxs, lens, bgs)
    np.savez_compressed(OUT / "concat_non_center.npz", feats=feats, nidxs=nidxs, lens=lens, bgs=bgs.astype(np.int32),
                        fwd=fwd, dfwd=dfwd, bwd=bwd)

    # ---- a15 permutate (test_permutate.py:8-13) ---------------------------------------------------
    pn, fd = 50, 11
    o2p = rng.permutation(pn).astype(np.int32)
    p2o = np.empty_like(o2p)
    p2o[o2p] = np.arange(pn, dtype=np.int32)
    feats = rng.uniform(-1, 1, (pn, fd)).astype(np.float32)
    np.savez_compressed(OUT / "permutate.npz", feats=feats, o2p=o2p, p2o=p2o,
                        fwd=R["permutate_forward"](feats, o2p), back=R["permutate_forward"](feats, p2o))
    print("wrote", sorted(p.name for p in OUT.glob("*.npz")))
    return 0


if __name__ == "__main__":
    sys.exit(main())

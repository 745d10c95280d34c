"""CPU: pin the oracle (oracle/pcs_oracle.c) against golden vectors produced by executing the
reference's own numpy test oracles (tests/golden/make_golden.py).  Tolerances are the reference
tests' own: mean|d| <= 1e-5 and max|d| <= 1e-4 (tf_ops/test/test_neighbor_scatter.py:194-199);
copies and index outputs must be bit-exact."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import pcs_oracle as orc


def close(a, b):
    d = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
    assert a.shape == b.shape
    assert d.size == 0 or (d.mean() <= 1e-5 and d.max() <= 1e-4), (d.mean(), d.max())


@pytest.mark.parametrize("tag", ["small", "wide", "odd"])
def test_neighbor_scatter_gather(tag):
    g = np.load(GOLDEN / f"neighbor_scatter_{tag}.npz")
    a = (g["nidxs"], g["lens"], g["bgs"])
    assert np.array_equal(orc.neighbor_scatter(g["feats"], *a, use_diff=False), g["copy_fwd"].astype(np.float32))
    close(orc.neighbor_scatter(g["feats"], *a, use_diff=True), g["diff_fwd"])
    close(orc.neighbor_gather(g["dsfeats"], *a, use_diff=False), g["copy_bwd"])
    close(orc.neighbor_gather(g["dsfeats"], *a, use_diff=True), g["diff_bwd"])


@pytest.mark.parametrize("tag", ["small", "wide"])
def test_sum_scatter_gather(tag):
    g = np.load(GOLDEN / f"sum_scatter_{tag}.npz")
    assert np.array_equal(orc.row_gather(g["feats"], g["cens"]), g["fwd"].astype(np.float32))
    close(orc.segment_sum(g["dsfeats"], g["lens"], g["bgs"]), g["bwd"])


@pytest.mark.parametrize("tag", ["pool", "ragged"])
def test_max_pool_unpool(tag):
    g = np.load(GOLDEN / f"max_pool_{tag}.npz")
    pf, mi = orc.max_feat_gather(g["ifeats"], g["vlens"], g["vbegs"])
    assert np.array_equal(pf, g["pfeats"].astype(np.float32))
    assert np.array_equal(mi, g["max_idxs"])  # np.argmax = first maximum = strict '<'
    d = orc.max_feat_scatter(g["dpfeats"], mi, g["vbegs"], g["vlens"], g["ifeats"].shape[0])
    assert np.array_equal(d, g["difeats"].astype(np.float32))
    assert np.array_equal(orc.row_gather(g["up_in"], g["vcens"]), g["up"].astype(np.float32))
    close(orc.segment_sum(g["dup"], g["vlens"], g["vbegs"]), g["dup_in"])


def test_concat_non_center():
    """eliminate_center + sum_feat_scatter + neighbor_scatter == np_concat_non_center_forward
    (tf_ops/graph_conv_layer.py:297-302 vs tf_ops/test/np_ops.py:79-114)."""
    g = np.load(GOLDEN / "concat_non_center.npz")
    on, ol, ob, oc = orc.eliminate_center(g["nidxs"], g["lens"], g["bgs"])
    assert np.array_equal(ol, g["lens"] - 1)
    assert np.array_equal(ob, g["bgs"] - np.arange(len(ol)))
    f1 = orc.row_gather(g["feats"], oc)
    f2 = orc.neighbor_scatter(g["feats"], on, ol, ob, use_diff=False)
    assert np.array_equal(np.concatenate([f1, f2], 1), g["fwd"].astype(np.float32))
    fd = g["feats"].shape[1]
    d1 = orc.segment_sum(g["dfwd"][:, :fd], ol, ob)
    d2 = orc.neighbor_gather(g["dfwd"][:, fd:], on, ol, ob, use_diff=False)
    close(d1 + d2, g["bwd"])


def test_permutate():
    g = np.load(GOLDEN / "permutate.npz")
    assert np.array_equal(orc.row_gather(g["feats"], g["o2p"]), g["fwd"])
    assert np.array_equal(orc.row_gather(g["feats"], g["p2o"]), g["back"])
    assert np.array_equal(orc.row_gather(orc.row_gather(g["feats"], g["o2p"]), g["p2o"]), g["feats"])

"""TEST INFRASTRUCTURE, NOT PRODUCT.  numpy restatement of the pooling composites of the reference's
Python op layer (tf_ops/graph_pooling_layer.py:39-352) over the C oracle (oracle/pcs_oracle.py):
what `points_pooling`, `average_downsample`, `context_points_pooling`, `class_pooling`,
`points_pooling_two_layers`, `context_points_pooling_two_layers`, `points_pooling_three_layers` and
`voxel_downsample` return for one block, in the reference's own arithmetic.

Arithmetic: a voxel centroid is `neighbor_sum_feat_gather(pts) / float32(vlens)` (graph_pooling_layer.py:57-58);
the reference kernel behind it accumulates `+=` in fp32 in edge order (sumFeatGather,
TFNeighborKernelNew.cu:19-25), so the centroid here is the sequential fp32 sum divided in fp32 — every
output of a composite is then a bit-exact function of the inputs, indices AND coordinates.

Written level-wise (one `_Level` per voxelisation, a `_reorder` per repermutation) rather than as the
reference's straight-line scripts; the per-function docstrings give the lines they restate.
"""
from __future__ import annotations

import numpy as np

from . import pcs_oracle as orc


def centroid(pts, vlens, vbegs):
    """graph_pooling_layer.py:57-58 / :124-125 — fp32 sum in edge order, fp32 divide."""
    s = orc.segment_sum_f32(pts, vlens, vbegs)
    return (s / vlens.astype(np.float32)[:, None]).astype(np.float32)


class _Level:
    """One voxelisation of a point set: voxel index -> permutation info -> permuted points ->
    centroids -> offsets (the block every composite repeats, e.g. graph_pooling_layer.py:41-60)."""

    def __init__(self, pts, voxel, block, classes=None):
        pts = np.ascontiguousarray(pts, np.float32)
        self.o2p, self.p2o, self.lens, self.begs, self.cens = orc.permutation_info(
            orc.voxel_index(pts, voxel, block), classes)
        self.pts = orc.row_gather(pts, self.o2p)
        self.centres = centroid(self.pts, self.lens, self.begs)
        self.dpts = orc.diff_xyz(self.pts, self.centres, self.cens)

    def carry(self, a):
        """PermutateFeature of a per-point array ([pn] or [pn,k], 4-byte dtype) by this level's o2p."""
        a = np.ascontiguousarray(a)
        return orc.row_gather(a[:, None], self.o2p)[:, 0] if a.ndim == 1 else orc.row_gather(a, self.o2p)


def _reorder(fine, o2p_coarse):
    """ComputeRepermutationInfo of `fine` (its voxels = the coarse level's points) after the coarse
    level's points were permuted by `o2p_coarse` (graph_pooling_layer.py:132-138).  Mutates `fine`'s
    CSR and point arrays; returns the row permutation for the arrays the caller carries."""
    ro2p, _, fine.lens, fine.begs, fine.cens = orc.repermutation_info(o2p_coarse, fine.lens, fine.begs,
                                                                      fine.pts.shape[0])
    fine.pts = orc.row_gather(fine.pts, ro2p)
    fine.dpts = orc.row_gather(fine.dpts, ro2p)
    return ro2p


def _rows(a, perm):
    a = np.ascontiguousarray(a)
    return orc.row_gather(a[:, None], perm)[:, 0] if a.ndim == 1 else orc.row_gather(a, perm)


def _i32(a):
    return np.ascontiguousarray(a).astype(np.int32)


def points_pooling(xyzs, feats, labels, voxel_size=0.2, block_size=3.0):
    """graph_pooling_layer.py:39-65."""
    lv = _Level(xyzs, voxel_size, block_size)
    return (lv.pts, lv.centres, lv.dpts, lv.carry(np.float32(feats)), lv.carry(_i32(labels)).astype(np.int64),
            lv.lens, lv.begs, lv.cens)


def average_downsample(xyzs, feats, ds_size=0.2, min_coordinate=3.0):
    """graph_pooling_layer.py:68-86 — block_size = 2*min_coordinate; centroids of points and features."""
    lv = _Level(xyzs, ds_size, min_coordinate * 2)
    return lv.centres, centroid(lv.carry(np.float32(feats)), lv.lens, lv.begs)


def context_points_pooling(xyzs, feats, ctx_idxs, voxel_size=0.2, block_size=3.0):
    """graph_pooling_layer.py:88-108 — ctx_idxs (indices INTO the points) follow through p2o (:98)."""
    lv = _Level(xyzs, voxel_size, block_size)
    return (lv.pts, lv.centres, lv.dpts, lv.carry(np.float32(feats)), lv.lens, lv.begs, lv.cens,
            lv.p2o[np.asarray(ctx_idxs, np.int64)])


def class_pooling(xyzs, feats, classes, labels, voxel_size=0.2, block_size=3.0):
    """graph_pooling_layer.py:230-257 — the class id is the least significant key field."""
    lv = _Level(xyzs, voxel_size, block_size, _i32(classes))
    return (lv.pts, lv.centres, lv.dpts, lv.carry(np.float32(feats)), lv.carry(_i32(labels)).astype(np.int64),
            lv.lens, lv.begs, lv.cens)


def voxel_downsample(xyzs, feats, labels, voxel_size, block_size):
    """graph_pooling_layer.py:330-352 (upstream's return line names undefined variables; these are the
    values computed under the names it evidently meant — labels stay int32, no cast, :341-343)."""
    lv = _Level(xyzs, voxel_size, block_size)
    return (lv.pts, lv.centres, lv.dpts, lv.carry(np.float32(feats)), lv.carry(_i32(labels)), lv.lens, lv.begs,
            lv.cens)


def _hierarchy(xyzs, voxels, block_size, carried):
    """N-level pooling: levels[k] voxelises the centroids of levels[k-1]; after each new level every
    finer level is repermuted so that its voxel runs stay contiguous, finest last
    (two levels: graph_pooling_layer.py:110-146; three: :261-325)."""
    levels = [_Level(xyzs, voxels[0], block_size)]
    carried = [levels[0].carry(c) for c in carried]
    for k in range(1, len(voxels)):
        levels.append(_Level(levels[-1].centres, voxels[k], block_size))
        perm = levels[k].o2p
        for fine in reversed(levels[:k]):  # level k-1 first, then k-2 with the permutation just made
            perm = _reorder(fine, perm)
        carried = [_rows(c, perm) for c in carried]
    xs = [lv.pts for lv in levels] + [levels[-1].centres]
    return (xs, [lv.dpts for lv in levels], carried, [lv.lens for lv in levels], [lv.begs for lv in levels],
            [lv.cens for lv in levels])


def points_pooling_two_layers(xyzs, feats, labels, voxel_size1, voxel_size2, block_size=3.0):
    """graph_pooling_layer.py:110-146."""
    xs, ds, (f, lab), vl, vb, vc = _hierarchy(xyzs, [voxel_size1, voxel_size2], block_size,
                                              [np.float32(feats), _i32(labels)])
    return xs, ds, f, lab.astype(np.int64), vl, vb, vc


def context_points_pooling_two_layers(xyzs, feats, labels, ctx_idxs, voxel_size1, voxel_size2, block_size=3.0):
    """graph_pooling_layer.py:149-189 — ctx_idxs is carried as a per-point int column (:151,:160,:185)."""
    xs, ds, (f, lab, ctx), vl, vb, vc = _hierarchy(xyzs, [voxel_size1, voxel_size2], block_size,
                                                   [np.float32(feats), _i32(labels), _i32(ctx_idxs)])
    return xs, ds, f, lab.astype(np.int64), vl, vb, vc, ctx


def points_pooling_three_layers(xyzs, feats, labels, voxel_size1, voxel_size2, voxel_size3, block_size=3.0):
    """graph_pooling_layer.py:261-325 (second repermutation chain :307-319)."""
    xs, ds, (f, lab), vl, vb, vc = _hierarchy(xyzs, [voxel_size1, voxel_size2, voxel_size3], block_size,
                                              [np.float32(feats), _i32(labels)])
    return xs, ds, f, lab.astype(np.int64), vl, vb, vc


# ---- the reference's structural checks, vectorised -------------------------------------------------
def check_vidxs(lens, begs, cens, n_fine=None):
    """PoolingIndex/test_permutation.py:55-66 + model_pooling.py:774-780: begs is the exclusive scan of
    lens, cens[begs[i]+j] == i, every point belongs to a voxel, the last run ends at the point count."""
    lens, begs, cens = (np.asarray(a, np.int64) for a in (lens, begs, cens))
    assert begs[0] == 0 and np.array_equal(begs[1:], np.cumsum(lens)[:-1])
    assert np.array_equal(cens, np.repeat(np.arange(len(lens)), lens))
    assert cens.max() < len(lens) and begs[-1] + lens[-1] == len(cens)
    if n_fine is not None:
        assert len(cens) == n_fine


def check_dxyzs(pts1, pts2, dpts1, vcens, tol=1e-6):
    """PoolingIndex/test_permutation.py:78-85 (prints there; asserted here): offset + centre == point."""
    err = np.abs(dpts1.astype(np.float64) + pts2[vcens].astype(np.float64) - pts1.astype(np.float64))
    assert err.max() <= tol, err.max()

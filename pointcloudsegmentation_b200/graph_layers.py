"""Layer-level wrappers over neighbor_ops / pooling_ops: the thin L3 functions every reference network
calls (tf_ops/graph_conv_layer.py:277-312,788-792 and tf_ops/graph_pooling_layer.py:20-352), same
names, same argument order, same return tuples.  Only the lines that touch the custom ops are
restated; the `fully_connected` glue around them is the caller's (plain PyTorch).

Extension for B200: every pooling / search wrapper takes an optional block description
(`block_ids` per point, or `block_begs` offsets) so that many independent room blocks can ride in one
block-diagonal launch — one 4-10k-point block alone cannot fill 148 SMs.  Without it the behaviour is
exactly the reference's single-block one.
"""
from __future__ import annotations

import torch

from . import neighbor_ops, pooling_ops


# ---- tf_ops/graph_conv_layer.py:277-312 ------------------------------------------------------------
def graph_pool(feats, vlens, vlens_bgs):
    pool_feats, _ = neighbor_ops.neighbor_max_feat_gather(feats, vlens, vlens_bgs)
    return pool_feats


def graph_pool_for_test(feats, vlens, vlens_bgs):
    return neighbor_ops.neighbor_max_feat_gather(feats, vlens, vlens_bgs)


def graph_avg_pool(feats, vlens, vbegs, vcens):
    feats = neighbor_ops.neighbor_sum_feat_gather(feats, vcens, vlens, vbegs)
    return feats / vlens.to(torch.float32).unsqueeze(1)


def graph_unpool(feats, vlens, vlens_bgs, cidxs):
    return neighbor_ops.neighbor_sum_feat_scatter(feats, cidxs, vlens, vlens_bgs)


def graph_eliminate_center(nidxs, nidxs_lens, nidxs_bgs):
    return neighbor_ops.eliminate_center(nidxs, nidxs_lens, nidxs_bgs)


def graph_concat_non_center_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs):
    nc_nidxs, nc_lens, nc_bgs, nc_cidxs = neighbor_ops.eliminate_center(nidxs, nidxs_lens, nidxs_bgs)
    sfeats1 = neighbor_ops.neighbor_sum_feat_scatter(ifeats, nc_cidxs, nc_lens, nc_bgs)
    sfeats2 = neighbor_ops.neighbor_scatter(ifeats, nc_nidxs, nc_lens, nc_bgs, use_diff=False)
    return torch.cat([sfeats1, sfeats2], dim=1)


def graph_neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs):
    return neighbor_ops.neighbor_scatter(ifeats, nidxs, nidxs_lens, nidxs_bgs, use_diff=False)


def graph_neighbor_sum(ifeats, nidxs_lens, nidxs_bgs, cidxs):
    return neighbor_ops.neighbor_sum_feat_gather(ifeats, cidxs, nidxs_lens, nidxs_bgs)


def graph_concat_scatter(feats, nidxs, nlens, nbegs, ncens):
    # tf_ops/graph_conv_layer.py:788-792
    scatter_feats1 = graph_unpool(feats, nlens, nbegs, ncens)
    scatter_feats2 = graph_neighbor_scatter(feats, nidxs, nlens, nbegs)
    return torch.cat([scatter_feats1, scatter_feats2], dim=1)


# ---- tf_ops/graph_pooling_layer.py:20-36 ------------------------------------------------------------
# (the layers go through the one-traversal form of csrc/search_multi.cu with a single graph: the same rows as the ops
# pooling_ops.search_neighborhood_brute_force[_range], tests/test_gpu_pooling_ops.py, for one candidate pass instead of two)
def search_neighborhood(xyzs, radius, block_begs=None, max_block_pn=None):
    return search_neighborhood_multi(xyzs, [(None, radius)], block_begs=block_begs, max_block_pn=max_block_pn)[0]


def search_neighborhood_range(xyzs, min_radius, max_radius, block_begs=None, max_block_pn=None):
    return search_neighborhood_multi(xyzs, [(min_radius, max_radius)], block_begs=block_begs, max_block_pn=max_block_pn)[0]


def search_neighborhood_multi(xyzs, radii, block_begs=None, max_block_pn=None):
    """[(min_radius or None, max_radius), ...] -> [search_neighborhood / search_neighborhood_range result, ...] from one
    traversal of the points (extension; the towers build up to four graphs per level)."""
    from . import ops

    specs = [(None if lo is None else lo * lo, hi * hi) for lo, hi in radii]
    return ops.search_neighborhood_multi(xyzs, specs, block_begs=block_begs, max_block_pn=max_block_pn)


def search_neighborhood_fixed(xyzs, radius, fixed_size=10, block_begs=None, max_block_pn=None):
    return pooling_ops.search_neighborhood_fixed_brute_force(
        xyzs, squared_nn_size=radius * radius, fixed_size=fixed_size, block_begs=block_begs,
        max_block_pn=max_block_pn)


def search_neighborhood_fixed_range(xyzs, min_radius, max_radius, fixed_size, block_begs=None,
                                    max_block_pn=None):
    return pooling_ops.search_neighborhood_fixed_brute_force_range(
        xyzs, squared_min_nn_size=min_radius * min_radius, squared_max_nn_size=max_radius * max_radius,
        fixed_size=fixed_size, block_begs=block_begs, max_block_pn=max_block_pn)


# ---- pooling composites (tf_ops/graph_pooling_layer.py:39-352) ----------------------------------------
def _avg(pts, vcens, vlens, vbegs):
    c = neighbor_ops.neighbor_sum_feat_gather(pts, vcens, vlens, vbegs)
    return c / vlens.to(torch.float32).unsqueeze(1)


def _voxelize(xyzs, voxel_size, block_size, block_ids=None):
    vidxs = pooling_ops.compute_voxel_index(xyzs, voxel_len=voxel_size, block_size=block_size)
    return pooling_ops.compute_permutation_info(vidxs, block_ids=block_ids)


def _labels_in(labels):
    return labels.to(torch.int32).unsqueeze(1)


def _labels_out(labels):
    return labels.squeeze(1).to(torch.int64)


def block_begs_from_ids(sorted_block_ids, num_blocks):
    """int32[nb+1] offsets from a non-decreasing per-point block id vector (batch extension)."""
    counts = torch.bincount(sorted_block_ids.to(torch.int64), minlength=num_blocks)
    begs = torch.zeros(num_blocks + 1, dtype=torch.int32, device=sorted_block_ids.device)
    begs[1:] = torch.cumsum(counts, 0).to(torch.int32)
    return begs


def points_pooling(xyzs, feats, labels, voxel_size=0.2, block_size=3.0):
    # tf_ops/graph_pooling_layer.py:39-65
    o2p, p2o, vlens, vbegs, vcens = _voxelize(xyzs, voxel_size, block_size)
    ppts = pooling_ops.permutate_feature(xyzs, o2p, p2o)
    pfeats = pooling_ops.permutate_feature(feats, o2p, p2o)
    plabels = _labels_out(pooling_ops.permutate_feature(_labels_in(labels), o2p, p2o))
    cpts = _avg(ppts, vcens, vlens, vbegs)
    dpts = pooling_ops.compute_diff_xyz(ppts, cpts, vcens)
    return ppts, cpts, dpts, pfeats, plabels, vlens, vbegs, vcens


def average_downsample(xyzs, feats, ds_size=0.2, min_coordinate=3.0):
    # tf_ops/graph_pooling_layer.py:68-86
    o2p, p2o, vlens, vbegs, vcens = _voxelize(xyzs, ds_size, min_coordinate * 2)
    ppts = pooling_ops.permutate_feature(xyzs, o2p, p2o)
    pfeats = pooling_ops.permutate_feature(feats, o2p, p2o)
    return _avg(ppts, vcens, vlens, vbegs), _avg(pfeats, vcens, vlens, vbegs)


def context_points_pooling(xyzs, feats, ctx_idxs, voxel_size=0.2, block_size=3.0):
    # tf_ops/graph_pooling_layer.py:88-108
    o2p, p2o, vlens, vbegs, vcens = _voxelize(xyzs, voxel_size, block_size)
    ppts = pooling_ops.permutate_feature(xyzs, o2p, p2o)
    pfeats = pooling_ops.permutate_feature(feats, o2p, p2o)
    ctx_idxs = p2o[ctx_idxs.to(torch.int64)]
    cpts = _avg(ppts, vcens, vlens, vbegs)
    dpts = pooling_ops.compute_diff_xyz(ppts, cpts, vcens)
    return ppts, cpts, dpts, pfeats, vlens, vbegs, vcens, ctx_idxs


def class_pooling(xyzs, feats, classes, labels, voxel_size=0.2, block_size=3.0):
    # tf_ops/graph_pooling_layer.py:230-257
    vidxs = pooling_ops.compute_voxel_index(xyzs, voxel_len=voxel_size, block_size=block_size)
    o2p, p2o, vlens, vbegs, vcens = pooling_ops.compute_permutation_info_with_class(vidxs, classes.to(torch.int32))
    ppts = pooling_ops.permutate_feature(xyzs, o2p, p2o)
    pfeats = pooling_ops.permutate_feature(feats, o2p, p2o)
    plabels = _labels_out(pooling_ops.permutate_feature(_labels_in(labels), o2p, p2o))
    cpts = _avg(ppts, vcens, vlens, vbegs)
    dpts = pooling_ops.compute_diff_xyz(ppts, cpts, vcens)
    return ppts, cpts, dpts, pfeats, plabels, vlens, vbegs, vcens


def _two_layers(xyzs, feats, labels, extra, voxel_size1, voxel_size2, block_size, block_ids):
    """Shared body of points_pooling_two_layers / context_points_pooling_two_layers
    (tf_ops/graph_pooling_layer.py:110-146,149-189).  `extra`: more [pn,k] int32 columns to carry."""
    # permutation 1
    o2p1, p2o1, vlens1, vbegs1, vcens1 = _voxelize(xyzs, voxel_size1, block_size, block_ids)
    pts1 = pooling_ops.permutate_feature(xyzs, o2p1, p2o1)
    feats = pooling_ops.permutate_feature(feats, o2p1, p2o1)
    labels = pooling_ops.permutate_feature(labels, o2p1, p2o1)
    extra = [pooling_ops.permutate_feature(e, o2p1, p2o1) for e in extra]
    bids1 = bids2 = None
    if block_ids is not None:
        bids1 = pooling_ops.permutate_feature(block_ids.unsqueeze(1), o2p1, p2o1).squeeze(1)
        bids2 = bids1[vbegs1.to(torch.int64)].contiguous()
    # centre 1
    pts2 = _avg(pts1, vcens1, vlens1, vbegs1)
    dpts1 = pooling_ops.compute_diff_xyz(pts1, pts2, vcens1)
    # permutation 2
    o2p2, p2o2, vlens2, vbegs2, vcens2 = _voxelize(pts2, voxel_size2, block_size, bids2)
    pts2 = pooling_ops.permutate_feature(pts2, o2p2, p2o2)
    bids3 = None
    if block_ids is not None:
        bids2 = pooling_ops.permutate_feature(bids2.unsqueeze(1), o2p2, p2o2).squeeze(1)
        bids3 = bids2[vbegs2.to(torch.int64)].contiguous()
    # centre 2
    pts3 = _avg(pts2, vcens2, vlens2, vbegs2)
    dpts2 = pooling_ops.compute_diff_xyz(pts2, pts3, vcens2)
    # repermutate 1
    ro2p1, rp2o1, vlens1, vbegs1, vcens1 = pooling_ops.compute_repermutation_info(o2p2, vlens1, vbegs1, vcens1)
    pts1 = pooling_ops.permutate_feature(pts1, ro2p1, rp2o1)
    dpts1 = pooling_ops.permutate_feature(dpts1, ro2p1, rp2o1)
    feats = pooling_ops.permutate_feature(feats, ro2p1, rp2o1)
    labels = pooling_ops.permutate_feature(labels, ro2p1, rp2o1)
    extra = [pooling_ops.permutate_feature(e, ro2p1, rp2o1) for e in extra]
    if block_ids is not None:
        bids1 = pooling_ops.permutate_feature(bids1.unsqueeze(1), ro2p1, rp2o1).squeeze(1)
    return ([pts1, pts2, pts3], [dpts1, dpts2], feats, labels, [vlens1, vlens2], [vbegs1, vbegs2],
            [vcens1, vcens2], extra, [bids1, bids2, bids3])


def points_pooling_two_layers(xyzs, feats, labels, voxel_size1, voxel_size2, block_size=3.0, block_ids=None):
    """tf_ops/graph_pooling_layer.py:110-146.  With `block_ids` (int32[pn], extension) an 8th element
    is appended to the returned tuple: the per-level sorted block ids [bids1, bids2, bids3]."""
    xs, dxs, feats, labels, vl, vb, vc, _, bids = _two_layers(
        xyzs, feats, _labels_in(labels), [], voxel_size1, voxel_size2, block_size, block_ids)
    out = (xs, dxs, feats, _labels_out(labels), vl, vb, vc)
    return out if block_ids is None else out + (bids,)


def context_points_pooling_two_layers(xyzs, feats, labels, ctx_idxs, voxel_size1, voxel_size2, block_size=3.0):
    # tf_ops/graph_pooling_layer.py:149-189
    xs, dxs, feats, labels, vl, vb, vc, extra, _ = _two_layers(
        xyzs, feats, _labels_in(labels), [ctx_idxs.to(torch.int32).unsqueeze(1)], voxel_size1, voxel_size2,
        block_size, None)
    return xs, dxs, feats, _labels_out(labels), vl, vb, vc, extra[0].squeeze(1)


def points_pooling_three_layers(xyzs, feats, labels, voxel_size1, voxel_size2, voxel_size3, block_size=3.0):
    # tf_ops/graph_pooling_layer.py:261-325
    xs, dxs, feats, labels, vl, vb, vc, _, _ = _two_layers(
        xyzs, feats, _labels_in(labels), [], voxel_size1, voxel_size2, block_size, None)
    pts1, pts2, pts3 = xs
    dpts1, dpts2 = dxs
    (vlens1, vlens2), (vbegs1, vbegs2), (vcens1, vcens2) = vl, vb, vc
    # permutation 3
    o2p3, p2o3, vlens3, vbegs3, vcens3 = _voxelize(pts3, voxel_size3, block_size)
    pts3 = pooling_ops.permutate_feature(pts3, o2p3, p2o3)
    pts4 = _avg(pts3, vcens3, vlens3, vbegs3)
    dpts3 = pooling_ops.compute_diff_xyz(pts3, pts4, vcens3)
    # repermutate 2
    ro2p2, rp2o2, vlens2, vbegs2, vcens2 = pooling_ops.compute_repermutation_info(o2p3, vlens2, vbegs2, vcens2)
    pts2 = pooling_ops.permutate_feature(pts2, ro2p2, rp2o2)
    dpts2 = pooling_ops.permutate_feature(dpts2, ro2p2, rp2o2)
    # repermutate 1
    ro2p1, rp2o1, vlens1, vbegs1, vcens1 = pooling_ops.compute_repermutation_info(ro2p2, vlens1, vbegs1, vcens1)
    pts1 = pooling_ops.permutate_feature(pts1, ro2p1, rp2o1)
    dpts1 = pooling_ops.permutate_feature(dpts1, ro2p1, rp2o1)
    feats = pooling_ops.permutate_feature(feats, ro2p1, rp2o1)
    labels = pooling_ops.permutate_feature(labels, ro2p1, rp2o1)
    return ([pts1, pts2, pts3, pts4], [dpts1, dpts2, dpts3], feats, _labels_out(labels),
            [vlens1, vlens2, vlens3], [vbegs1, vbegs2, vbegs3], [vcens1, vcens2, vcens3])


def permutate_feats(feats, o2p, p2o):
    return pooling_ops.permutate_feature(feats, o2p, p2o)


def voxel_downsample(xyzs, feats, labels, voxel_size, block_size):
    """tf_ops/graph_pooling_layer.py:330-352.  (Upstream returns undefined names — a NameError at
    call time; this returns the values those names evidently meant.)"""
    o2p, p2o, vlens, vbegs, vcens = _voxelize(xyzs, voxel_size, block_size)
    xyzs = pooling_ops.permutate_feature(xyzs, o2p, p2o)
    feats = pooling_ops.permutate_feature(feats, o2p, p2o)
    labels = pooling_ops.permutate_feature(labels.to(torch.int32).unsqueeze(1), o2p, p2o).squeeze(1)
    ds_xyzs = _avg(xyzs, vcens, vlens, vbegs)
    dxyzs = pooling_ops.compute_diff_xyz(xyzs, ds_xyzs, vcens)
    return xyzs, ds_xyzs, dxyzs, feats, labels, vlens, vbegs, vcens

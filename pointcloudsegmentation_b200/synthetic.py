"""Seeded synthetic point clouds of the reference's dataset shapes (no datasets on the box).

S3DIS block geometry from s3dis_util.py:12-18: 3 m x 3 m blocks, sample stride 0.05 m, between 512 and
10240 points, block-normalised so xy in [-1.5, 1.5] and z >= 0 (s3dis_util.py:118-120).  At stride 0.05 m
a surface carries ~400 points/m^2, which is what gives the mean degrees the reference notes in its
comments (r=0.1 -> 12, r=0.15 -> 29: model_pointnet.py:935,953).
"""
from __future__ import annotations

import numpy as np


def s3dis_block(rng: np.random.Generator, pn: int = 8192, block: float = 3.0, height: float = 2.7,
                stride: float = 0.05) -> np.ndarray:
    """Room corner: floor + two adjacent walls on a jittered `stride` lattice, randomly thinned to
    `pn` points, 2 % of them replaced by volume clutter.  Returns float32 [pn,3]."""
    h = block / 2
    g = np.arange(-h + stride / 2, h, stride)
    gz = np.arange(stride / 2, height, stride)
    fx, fy = np.meshgrid(g, g, indexing="ij")
    floor = np.stack([fx.ravel(), fy.ravel(), np.zeros(fx.size)], 1)
    wy, wz = np.meshgrid(g, gz, indexing="ij")
    wall1 = np.stack([np.full(wy.size, -h), wy.ravel(), wz.ravel()], 1)
    wall2 = np.stack([wy.ravel(), np.full(wy.size, h), wz.ravel()], 1)
    pts = np.concatenate([floor, wall1, wall2])
    pts = pts + rng.uniform(-stride / 2, stride / 2, pts.shape) * np.array([1.0, 1.0, 1.0])
    # keep surfaces thin: squash the jitter along each surface normal
    n_f, n_w = floor.shape[0], wall1.shape[0]
    pts[:n_f, 2] = np.abs(rng.normal(0, 0.004, n_f))
    pts[n_f:n_f + n_w, 0] = -h + np.abs(rng.normal(0, 0.004, n_w))
    pts[n_f + n_w:, 1] = h - np.abs(rng.normal(0, 0.004, n_w))
    if pn > pts.shape[0]:
        raise ValueError(f"pn={pn} exceeds the {pts.shape[0]} lattice sites of this block")
    pts = pts[rng.choice(pts.shape[0], pn, replace=False)]
    n_noise = max(1, pn // 50)
    pts[:n_noise] = np.stack([rng.uniform(-h, h, n_noise), rng.uniform(-h, h, n_noise),
                              rng.uniform(0, height, n_noise)], 1)
    pts[:, :2] = np.clip(pts[:, :2], -h, h)
    pts[:, 2] = np.clip(pts[:, 2], 0, None)
    return pts[rng.permutation(pn)].astype(np.float32)


def s3dis_crop(rng: np.random.Generator, pn: int = 4096) -> np.ndarray:
    """`pn` points at FULL surface density (~400 pts/m^2, the stride-0.05 m sampling): a 10000-point
    block cropped to the pn points nearest the corner where floor and both walls meet.  This is the
    op-level config of BASELINE.json: 4096 points whose radius-0.1 graph has mean degree ~12
    (model_pointnet.py:953), where a random thinning of the block would give ~5."""
    full = s3dis_block(rng, 10000).astype(np.float64)
    corner = np.array([-1.5, 1.5, 0.0])
    d = np.abs(full - corner).max(axis=1)  # Chebyshev: keeps an axis-aligned sub-room
    keep = np.argsort(d, kind="stable")[:pn]
    return full[np.sort(keep)].astype(np.float32)


def s3dis_batch(seed: int, num_blocks: int, pn_min: int = 6144, pn_max: int = 10000, feat_dim: int = 3,
                num_classes: int = 13):
    """`num_blocks` blocks concatenated block-diagonally -> dict of numpy arrays:
    xyzs [P,3] f32, feats [P,feat_dim] f32 in [0,1] (rgb), labels [P] i64, block_ids [P] i32."""
    rng = np.random.default_rng(seed)
    xs, bs = [], []
    for b in range(num_blocks):
        pn = int(rng.integers(pn_min, pn_max + 1))
        xs.append(s3dis_block(rng, pn))
        bs.append(np.full(pn, b, np.int32))
    xyzs = np.concatenate(xs)
    P = xyzs.shape[0]
    return {"xyzs": xyzs, "feats": rng.uniform(0, 1, (P, feat_dim)).astype(np.float32),
            "labels": rng.integers(0, num_classes, P).astype(np.int64), "block_ids": np.concatenate(bs),
            "num_blocks": num_blocks}


def s3dis_batch_to_budget(seed: int, num_blocks: int, target_points: int, extra: int = 8, **kw):
    """s3dis_batch of `num_blocks` blocks chosen out of num_blocks + extra candidates so that the batch holds as close
    to `target_points` points as single swaps get (dist.fill_to_budget): what a loader does to give every data-parallel
    rank the same work per step.  Same dict as s3dis_batch."""
    from .dist import fill_to_budget

    d = s3dis_batch(seed, num_blocks + extra, **kw)
    sizes = np.bincount(d["block_ids"], minlength=num_blocks + extra)
    keep = fill_to_budget(sizes.tolist(), num_blocks, target_points)
    begs = np.concatenate([[0], np.cumsum(sizes)])
    sel = np.concatenate([np.arange(begs[b], begs[b + 1]) for b in keep])
    new_ids = np.repeat(np.arange(num_blocks, dtype=np.int32), sizes[keep])
    return {"xyzs": d["xyzs"][sel], "feats": d["feats"][sel], "labels": d["labels"][sel], "block_ids": new_ids,
            "num_blocks": num_blocks}


def modelnet_batch(seed: int, num_clouds: int = 32, pn: int = 1024):
    """ModelNet40-shaped batch (train_modelnet40: 1024 points per cloud, batch 32): each cloud lies on the
    surface of a random box or cylinder inscribed in the unit sphere.  -> xyzs [B*pn,3] f32, block_ids."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(num_clouds):
        if rng.random() < 0.5:  # box
            half = rng.uniform(0.2, 0.55, 3)
            face = rng.integers(0, 6, pn)
            p = rng.uniform(-1, 1, (pn, 3)) * half
            ax = face % 3
            p[np.arange(pn), ax] = np.where(face < 3, half[ax], -half[ax])
        else:  # cylinder (side + caps)
            r, h = rng.uniform(0.2, 0.6), rng.uniform(0.2, 0.6)
            th = rng.uniform(0, 2 * np.pi, pn)
            side = rng.random(pn) < 0.7
            rr = np.where(side, r, r * np.sqrt(rng.random(pn)))
            z = np.where(side, rng.uniform(-h, h, pn), np.where(rng.random(pn) < 0.5, h, -h))
            p = np.stack([rr * np.cos(th), rr * np.sin(th), z], 1)
        out.append(p)
    xyzs = np.concatenate(out).astype(np.float32)
    return {"xyzs": xyzs, "block_ids": np.repeat(np.arange(num_clouds, dtype=np.int32), pn)}


def semantic3d_block(seed: int, pn: int = 1_000_000, extent: float = 80.0):
    """Semantic3D-shaped dense block (train_gpn_semantic3d_dense.py:58: voxels 0.45 / 1.5, radii 0.2-0.4 on
    clouds pre-sampled at ~0.1 m): ~pn points of a terrestrial scan tile of extent x extent m — ground
    plane (~80 pts/m^2), two 8 m facades, vegetation blobs.  -> xyzs [pn,3] f32, x, y in [0, extent]."""
    rng = np.random.default_rng(seed)
    n_g, n_f = int(pn * 0.5), int(pn * 0.35)
    n_v = pn - n_g - n_f
    g = np.stack([rng.uniform(0, extent, n_g), rng.uniform(0, extent, n_g), rng.normal(0, 0.01, n_g)], 1)
    u = rng.uniform(0, extent, n_f)
    z = 8.0 * rng.random(n_f)
    side = rng.random(n_f) < 0.5
    off = np.abs(rng.normal(0, 0.01, n_f))
    f = np.stack([np.where(side, extent - off, u), np.where(side, u, extent - off), z], 1)
    nb = 400
    c = rng.uniform(2.0, extent - 2.0, (nb, 2))
    k = rng.integers(0, nb, n_v)
    v = np.concatenate([c[k] + rng.normal(0, 1.0, (n_v, 2)), np.abs(rng.normal(2.0, 1.0, (n_v, 1)))], 1)
    pts = np.concatenate([g, f, v])
    pts[:, :2] = np.clip(pts[:, :2], 0, extent)
    return pts[rng.permutation(pn)].astype(np.float32)

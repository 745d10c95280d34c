"""TEST INFRASTRUCTURE, NOT PRODUCT.  numpy restatement of csrc/relu_concat_dropout.cu: classifier_v3's
`dropout(concat([relu(fc), pfeats], axis=3), 0.7)` (model.py:783-793) with a counter-based mask.

The reference draws its mask from TensorFlow's generator (tf.nn.dropout); no two frameworks share a random
stream, so parity for this op is: (1) the generator is Philox4x32-10 (Salmon et al., "Parallel random numbers:
as easy as 1, 2, 3", SC'11) and is pinned here against the Random123 known-answer vectors
(tests/test_dropout_mask_oracle.py); (2) given the mask, values and gradients are tf.nn.dropout's
(kept x / keep_prob, dropped 0), bit for bit against the CUDA kernel; (3) the kept fraction is keep_prob.
"""
from __future__ import annotations

import numpy as np

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_LO = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """ctr [n,4] uint32, key (k0, k1) -> [n,4] uint32."""
    c = np.asarray(ctr, dtype=np.uint64).reshape(-1, 4)
    x0, x1, x2, x3 = (c[:, i].copy() for i in range(4))
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = _M0 * x0, _M1 * x2  # 32x32 -> 64-bit products
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & _LO, p1 >> np.uint64(32), p1 & _LO
        x0, x1, x2, x3 = hi1 ^ x1 ^ np.uint64(k0), lo1, hi0 ^ x3 ^ np.uint64(k1), lo0
        k0, k1 = (k0 + _W0) & 0xFFFFFFFF, (k1 + _W1) & 0xFFFFFFFF
    return np.stack([x0, x1, x2, x3], 1).astype(np.uint32)


def keep_mask(M, W, keep_prob, seed):
    """bool [M, W]: element e = row*W + col is word j = (e % 128) // 32 of Philox counter (32 (e // 128) + e % 32, 0)
    under key = seed (lo, hi); kept iff word < keep_prob * 2^32."""
    if keep_prob >= 1.0:
        return np.ones((M, W), dtype=bool)
    total = M * W
    e = np.arange(total, dtype=np.uint64)
    j = ((e % np.uint64(128)) // np.uint64(32)).astype(np.int64)
    c = (e // np.uint64(128)) * np.uint64(32) + e % np.uint64(32)
    ctr = np.zeros((total, 4), dtype=np.uint64)
    ctr[:, 0], ctr[:, 1] = c & _LO, c >> np.uint64(32)
    words = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
    t = float(np.float32(keep_prob)) * 4294967296.0
    thresh = 4294967295 if t >= 4294967295.0 else int(t)
    return (words[np.arange(total), j].astype(np.uint64) < np.uint64(thresh)).reshape(M, W)


def relu_concat_dropout_forward(a, b, keep_prob, seed):
    x = np.concatenate([np.maximum(a, np.float32(0)), b], 1).astype(np.float32)
    if keep_prob >= 1.0:
        return x
    scale = np.float32(1.0) / np.float32(keep_prob)
    return np.where(keep_mask(x.shape[0], x.shape[1], keep_prob, seed), x * scale, np.float32(0)).astype(np.float32)


def relu_concat_dropout_backward(g, a, keep_prob, seed):
    """(da, db) for g of the forward's output; a is the forward's first input (sign of relu)."""
    M, W = g.shape
    Na = a.shape[1]
    scale = np.float32(1.0) if keep_prob >= 1.0 else np.float32(1.0) / np.float32(keep_prob)
    gm = np.where(keep_mask(M, W, keep_prob, seed), g * scale, np.float32(0)).astype(np.float32)
    return np.where(a > 0, gm[:, :Na], np.float32(0)).astype(np.float32), gm[:, Na:].copy()

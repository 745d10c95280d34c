/*
 * pcs_oracle_index.c — TEST INFRASTRUCTURE, NOT PRODUCT (see pcs_oracle.c header).
 * CPU restatement of the tf_ops/PoolingIndex kernels: voxel index, permutation info, re-permutation,
 * diff xyz, brute-force radius search.  All outputs integer or single-rounding fp32 => bit-exact.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* a12 — computeVoxelIdxKernel PoolingIndex/ComputeVoxelIdx.cu:4-27.  The float->unsigned store
 * compiles to F2I.U32.FLOOR (saturating: negative/NaN -> 0, >= 2^32 -> 0xffffffff). */
static inline int32_t f2u_floor_sat(float q) {
    if (!(q >= 0.f)) return 0; /* negative or NaN */
    const float f = floorf(q);
    if (f >= 4294967296.f) return (int32_t)0xffffffffu;
    return (int32_t)(uint32_t)f;
}
void orc_voxel_index(const float *pts, int64_t pn, int64_t stride, float voxel_len, float min_x, float min_y,
                     float min_z, int32_t *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        const float x = pts[i * stride], y = pts[i * stride + 1], z = pts[i * stride + 2];
        out[i * 3 + 0] = f2u_floor_sat((x - min_x) / voxel_len);
        out[i * 3 + 1] = f2u_floor_sat((y - min_y) / voxel_len);
        out[i * 3 + 2] = f2u_floor_sat((z - min_z) / voxel_len);
    }
}

/* a13/a14 — computePermutationInfo[WithClasses]Impl PoolingIndex/ComputePermutationInfo.cu:91-229:
 * key as computeWholeIdxs[WithClass]Kernel :10-44, stable ascending sort on the signed 64-bit key
 * (thrust::sort_by_key on `long` = stable radix sort), inverse map :58-68, run lengths
 * (reduce_by_key :131), exclusive scan :144, computeCens :71-88.
 * block_ids (extension, nullable): most significant sort field.  Returns vn. */
typedef struct { int64_t blk; int64_t key; int32_t idx; } orc_kv;
static int orc_kv_cmp(const void *a, const void *b) {
    const orc_kv *x = (const orc_kv *)a, *y = (const orc_kv *)b;
    if (x->blk != y->blk) return x->blk < y->blk ? -1 : 1;
    if (x->key != y->key) return x->key < y->key ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}
int64_t orc_permutation_info(const int32_t *vidx, const int32_t *classes, const int32_t *block_ids, int64_t pn,
                             int32_t *o2p, int32_t *p2o, int32_t *lens, int32_t *begs, int32_t *cens) {
    if (pn == 0) return 0;
    orc_kv *kv = (orc_kv *)malloc(sizeof(orc_kv) * (size_t)pn);
    for (int64_t i = 0; i < pn; ++i) {
        int64_t k = vidx[i * 3];
        k <<= 16; k += vidx[i * 3 + 1];
        k <<= 16; k += vidx[i * 3 + 2];
        if (classes) { k <<= 16; k += classes[i]; }
        kv[i].blk = block_ids ? block_ids[i] : 0;
        kv[i].key = k;
        kv[i].idx = (int32_t)i;
    }
    qsort(kv, (size_t)pn, sizeof(orc_kv), orc_kv_cmp);
    int64_t vn = 0;
    for (int64_t p = 0; p < pn; ++p) {
        o2p[p] = kv[p].idx;
        p2o[kv[p].idx] = (int32_t)p;
        if (p == 0 || kv[p].key != kv[p - 1].key || kv[p].blk != kv[p - 1].blk) {
            begs[vn] = (int32_t)p;
            lens[vn] = 0;
            ++vn;
        }
        lens[vn - 1] += 1;
        cens[p] = (int32_t)(vn - 1);
    }
    free(kv);
    return vn;
}

/* a16 — computeDiffXYZKernel PoolingIndex/ComputeDiffXYZ.cu:4-19 */
void orc_diff_xyz(const float *xyzs, const float *cxyzs, const int32_t *cidxs, int64_t pn1, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn1; ++i)
        for (int k = 0; k < 3; ++k) out[i * 3 + k] = xyzs[i * 3 + k] - cxyzs[(int64_t)cidxs[i] * 3 + k];
}

/* a17 — computeRepermutationInfoImpl PoolingIndex/ComputeRepermutationInfo.cu:20-88 */
void orc_repermutation_info(const int32_t *o2p2, const int32_t *lens, const int32_t *begs, int64_t pn1,
                            int64_t pn2, int32_t *ro2p1, int32_t *rp2o1, int32_t *rlens, int32_t *rbegs,
                            int32_t *rcens) {
    int32_t acc = 0;
    for (int64_t i = 0; i < pn2; ++i) { rlens[i] = lens[o2p2[i]]; rbegs[i] = acc; acc += rlens[i]; }
    for (int64_t i = 0; i < pn2; ++i) {
        const int32_t obeg = begs[o2p2[i]];
        for (int32_t j = 0; j < rlens[i]; ++j) { ro2p1[rbegs[i] + j] = obeg + j; rcens[rbegs[i] + j] = (int32_t)i; }
    }
    for (int64_t p = 0; p < pn1; ++p) rp2o1[ro2p1[p]] = (int32_t)p;
}

/* a18/a19 predicate — SearchNeighborhood.cu:24 / :50-51 as nvcc -O2 contracts it for every one of
 * the six kernels (read off the SASS: FMUL dy*dy; FFMA dx,dx; FFMA dz,dz; FSETP). */
static inline int orc_hit(const float *t, const float *c, float sq_min, float sq_max, int use_min) {
    const float dx = t[0] - c[0], dy = t[1] - c[1], dz = t[2] - c[2];
    const float d2 = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
    return use_min ? (d2 < sq_max && d2 > sq_min) : (d2 < sq_max);
}
static void orc_block_of(const int32_t *block_begs, int64_t nb, int64_t i, int64_t pn, int64_t *bb, int64_t *be) {
    *bb = 0; *be = pn;
    if (!block_begs) return;
    int64_t lo = 0, hi = nb; /* last b with block_begs[b] <= i */
    while (hi - lo > 1) { int64_t mid = (lo + hi) / 2; if (block_begs[mid] <= i) lo = mid; else hi = mid; }
    *bb = block_begs[lo]; *be = block_begs[lo + 1];
}

/* countNeighborNum[Range]Kernel SearchNeighborhood.cu:5-55 + exclusive_scan :143; returns en */
int64_t orc_search_count(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, float sq_min,
                         float sq_max, int use_min, int32_t *lens, int32_t *begs) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < pn; ++i) {
        int64_t bb, be;
        orc_block_of(block_begs, nb, i, pn, &bb, &be);
        int32_t cnt = 0;
        for (int64_t j = bb; j < be; ++j) cnt += orc_hit(xyzs + j * 3, xyzs + i * 3, sq_min, sq_max, use_min);
        lens[i] = cnt;
    }
    int64_t acc = 0;
    for (int64_t i = 0; i < pn; ++i) { begs[i] = (int32_t)acc; acc += lens[i]; }
    return acc;
}
/* computeNeighborIdxs[Range]Kernel SearchNeighborhood.cu:57-123 */
void orc_search_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, float sq_min,
                     float sq_max, int use_min, const int32_t *begs, int32_t *idxs, int32_t *cens) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < pn; ++i) {
        int64_t bb, be;
        orc_block_of(block_begs, nb, i, pn, &bb, &be);
        int64_t w = begs[i];
        for (int64_t j = bb; j < be; ++j)
            if (orc_hit(xyzs + j * 3, xyzs + i * 3, sq_min, sq_max, use_min)) { idxs[w] = (int32_t)j; cens[w] = (int32_t)i; ++w; }
    }
}
/* a20 — computeNeighborIdxsFixed[Range]Kernel SearchNeighborhoodFixed.cu:6-102 + computeNeighborhoodStats :139-152 */
void orc_search_fixed(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, float sq_min,
                      float sq_max, int use_min, int64_t K, int32_t *idxs, int32_t *lens, int32_t *begs,
                      int32_t *cens) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < pn; ++i) {
        int64_t bb, be;
        orc_block_of(block_begs, nb, i, pn, &bb, &be);
        int64_t cnt = 0;
        int32_t last = (int32_t)i;
        for (int64_t j = bb; j < be && cnt < K; ++j)
            if (orc_hit(xyzs + j * 3, xyzs + i * 3, sq_min, sq_max, use_min)) {
                idxs[i * K + cnt] = (int32_t)j; cens[i * K + cnt] = (int32_t)i; last = (int32_t)j; ++cnt;
            }
        for (; cnt < K; ++cnt) { idxs[i * K + cnt] = last; cens[i * K + cnt] = (int32_t)i; }
        lens[i] = (int32_t)K;
        begs[i] = (int32_t)(i * K);
    }
}

/* ---- libPointUtil restatements (un-vendored dependency; semantics fixed in include/pcseg_ops.h) -------------------
 * radius search of queries into sources (findNeighborRadiusCPU(xyzs, q_idxs, r), s3dis_util.py:66), same predicate
 * as the self search above; returns en. */
int64_t orc_cross_search_count(const float *s, int64_t sn, const float *q, int64_t qn, float sq_min, float sq_max,
                               int use_min, int32_t *lens, int32_t *begs) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < qn; ++i) {
        int32_t cnt = 0;
        for (int64_t j = 0; j < sn; ++j) cnt += orc_hit(s + j * 3, q + i * 3, sq_min, sq_max, use_min);
        lens[i] = cnt;
    }
    int64_t acc = 0;
    for (int64_t i = 0; i < qn; ++i) { begs[i] = (int32_t)acc; acc += lens[i]; }
    return acc;
}
void orc_cross_search_fill(const float *s, int64_t sn, const float *q, int64_t qn, float sq_min, float sq_max,
                           int use_min, const int32_t *begs, int32_t *idxs, int32_t *cens) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < qn; ++i) {
        int64_t w = begs[i];
        for (int64_t j = 0; j < sn; ++j)
            if (orc_hit(s + j * 3, q + i * 3, sq_min, sq_max, use_min)) { idxs[w] = (int32_t)j; cens[w] = (int32_t)i; ++w; }
    }
}
/* exact K nearest sources of every query (findNeighborInAnotherCPU(sxyzs, qxyzs, K), interpolate.py:136): ascending
 * (squared distance, index), squared distance in fp32 with the search's contraction.  Insertion into a sorted K-list. */
void orc_knn(const float *s, int64_t sn, const float *q, int64_t qn, int64_t K, int32_t *idxs, float *dists) {
#pragma omp parallel for schedule(dynamic, 64)
    for (int64_t i = 0; i < qn; ++i) {
        float bd[32]; int32_t bj[32]; int n = 0;
        const float *c = q + i * 3;
        for (int64_t j = 0; j < sn; ++j) {
            const float *t = s + j * 3;
            const float dx = t[0] - c[0], dy = t[1] - c[1], dz = t[2] - c[2];
            const float d2 = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
            if (n == K && !(d2 < bd[K - 1])) continue; /* equal distance, larger index: stays out */
            int p = n < K ? n : (int)K - 1;
            while (p > 0 && bd[p - 1] > d2) { bd[p] = bd[p - 1]; bj[p] = bj[p - 1]; --p; }
            bd[p] = d2; bj[p] = (int32_t)j;
            if (n < K) ++n;
        }
        for (int k = 0; k < K; ++k) { idxs[i * K + k] = k < n ? bj[k] : -1; dists[i * K + k] = k < n ? bd[k] : INFINITY; }
    }
}

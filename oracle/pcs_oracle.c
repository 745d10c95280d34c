/*
 * pcs_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.  CPU restatement of the reference's tf_ops/ CUDA
 * kernels (liuyuan-pal/PointCloudSegmentation), one function per reference kernel, each citing
 * the reference file:line whose loop nest it follows.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this; the product path
 * (pointcloudsegmentation_b200/) never does.
 *
 * Parity pinning: the float ops are checked against golden vectors produced by executing the
 * reference's own numpy test oracles (tests/golden/make_golden.py runs functions taken from
 * /root/reference/tf_ops/test/ scripts); the index ops, which the reference's tests only check
 * structurally, are additionally checked on the GPU box against oracle/_ref/libref.so — the
 * reference's unmodified .cu files compiled for sm_100a (oracle/Makefile).
 *
 * Arithmetic: integer/index outputs and single-rounding float outputs (copies, one subtract, one
 * multiply) are bit-exact restatements in fp32; sums are accumulated in double (the reference's
 * tests compare against float64 numpy, tf_ops/test/test_neighbor_scatter.py:8-63).
 * Build: gcc -O3 -mavx2 -mfma -ffp-contract=off -fopenmp (no fast-math; fmaf only where written).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* a1 — featScatter / diffFeatScatter, TFNeighborKernel.cu:5-59 */
void orc_neighbor_scatter(const float *ifeats, const int32_t *nidxs, const int32_t *lens, const int32_t *bgs,
                          int64_t pn, int64_t C, int use_diff, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        const int64_t bg = bgs[i];
        for (int64_t j = 0; j < lens[i]; ++j) {
            const int64_t n = nidxs[bg + j];
            for (int64_t c = 0; c < C; ++c) {
                const float v = ifeats[n * C + c];
                out[(bg + j) * C + c] = use_diff ? v - ifeats[i * C + c] : v;
            }
        }
    }
}

/* a2 — featGather / diffFeatGather, TFNeighborKernel.cu:63-125 (atomicAdd there; order-free here
 * because the accumulator is double).  Parallel over channel stripes => no write races. */
void orc_neighbor_gather(const float *sfeats, const int32_t *nidxs, const int32_t *lens, const int32_t *bgs,
                         int64_t pn, int64_t C, int use_diff, double *out) {
    memset(out, 0, sizeof(double) * (size_t)(pn * C));
    int nt = orc_num_threads();
    if (nt > C) nt = (int)(C > 0 ? C : 1);
#pragma omp parallel for schedule(static) num_threads(nt)
    for (int t = 0; t < nt; ++t) {
        const int64_t c0 = C * t / nt, c1 = C * (t + 1) / nt;
        for (int64_t i = 0; i < pn; ++i) {
            const int64_t bg = bgs[i];
            for (int64_t j = 0; j < lens[i]; ++j) {
                const int64_t n = nidxs[bg + j];
                for (int64_t c = c0; c < c1; ++c) {
                    const double v = sfeats[(bg + j) * C + c];
                    out[n * C + c] += v;
                    if (use_diff) out[i * C + c] -= v;
                }
            }
        }
    }
}

/* a3 / a7 — sumFeatGather TFNeighborKernelNew.cu:4-27; locWSumForwardKernel TFNeighborKernel.cu:128-151 */
void orc_segment_sum(const float *feats, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                     double *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        double *o = out + i * C;
        for (int64_t c = 0; c < C; ++c) o[c] = 0.0;
        for (int64_t j = 0; j < lens[i]; ++j) {
            const float *f = feats + (bgs[i] + j) * C;
            for (int64_t c = 0; c < C; ++c) o[c] += (double)f[c];
        }
    }
}

/* a3 in the reference's OWN arithmetic: sumFeatGather (TFNeighborKernelNew.cu:19-25) accumulates `+=` in fp32
 * in edge order; this is what the pooling composites' voxel centroids are made of
 * (tf_ops/graph_pooling_layer.py:57-58), so index outputs downstream of a centroid are pinned with it. */
void orc_segment_sum_f32(const float *feats, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                         float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        float *o = out + i * C;
        for (int64_t c = 0; c < C; ++c) o[c] = 0.f;
        for (int64_t j = 0; j < lens[i]; ++j) {
            const float *f = feats + (bgs[i] + j) * C;
            for (int64_t c = 0; c < C; ++c) o[c] = o[c] + f[c];
        }
    }
}

/* a4 / a15 — sumFeatScatter TFNeighborKernelNew.cu:32-47; permutateFeatureKernel PermutateFeature.cu:5-19 */
void orc_row_gather(const float *in, const int32_t *idx, int64_t n_rows, int64_t C, float *out) {
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < n_rows; ++r) memcpy(out + r * C, in + (int64_t)idx[r] * C, sizeof(float) * (size_t)C);
}

/* a5 — maxFeatGather TFNeighborKernelNew.cu:83-118 (memset :168; empty rows: idx defined as 0) */
void orc_max_feat_gather(const float *feats, const int32_t *lens, const int32_t *bgs, int64_t pn2, int64_t C,
                         float *gfeats, int32_t *max_idxs) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn2; ++i) {
        float *g = gfeats + i * C;
        int32_t *mi = max_idxs + i * C;
        for (int64_t c = 0; c < C; ++c) { g[c] = 0.f; mi[c] = 0; }
        if (lens[i] == 0) continue;
        const float *f0 = feats + (int64_t)bgs[i] * C;
        for (int64_t c = 0; c < C; ++c) g[c] = f0[c];
        for (int64_t j = 1; j < lens[i]; ++j) {
            const float *f = f0 + j * C;
            for (int64_t c = 0; c < C; ++c)
                if (g[c] < f[c]) { g[c] = f[c]; mi[c] = (int32_t)j; }
        }
    }
}

/* a6 — maxFeatScatter TFNeighborKernelNew.cu:121-139 (memset :204) */
void orc_max_feat_scatter(const float *gfeats, const int32_t *max_idxs, const int32_t *bgs, const int32_t *lens,
                          int64_t pn1, int64_t pn2, int64_t C, float *out) {
    memset(out, 0, sizeof(float) * (size_t)(pn1 * C));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn2; ++i) {
        if (lens[i] == 0) continue;
        for (int64_t c = 0; c < C; ++c)
            out[((int64_t)bgs[i] + max_idxs[i * C + c]) * C + c] = gfeats[i * C + c];
    }
}

/* a8 — locWSumBackwardKernel TFNeighborKernel.cu:153-177 (memset :452) */
void orc_locw_sum_backward(const float *dlw_sum, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t m,
                           int64_t csum, float *dlw) {
    memset(dlw, 0, sizeof(float) * (size_t)(csum * m));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i)
        for (int64_t j = 0; j < lens[i]; ++j)
            memcpy(dlw + (bgs[i] + j) * m, dlw_sum + i * m, sizeof(float) * (size_t)m);
}

/* a9 — locationWeightFeatSumForwardKernel TFNeighborKernel.cu:179-211 */
void orc_locw_feat_sum(const float *lw, const float *tfeats, const int32_t *lens, const int32_t *bgs, int64_t pn,
                       int64_t m, int64_t ofn, double *out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        double *o = out + i * m * ofn;
        for (int64_t k = 0; k < m * ofn; ++k) o[k] = 0.0;
        for (int64_t j = 0; j < lens[i]; ++j) {
            const int64_t e = bgs[i] + j;
            for (int64_t a = 0; a < m; ++a) {
                const double w = lw[e * m + a];
                for (int64_t q = 0; q < ofn; ++q) o[a * ofn + q] += w * (double)tfeats[(e * m + a) * ofn + q];
            }
        }
    }
}

/* a10 — locationWeightFeatSumBackwardKernel TFNeighborKernel.cu:213-252.  dtfeats is one fp32
 * multiply (bit-exact); dlw is a sum over ofn (double). Edges outside every segment: dlw = 0
 * (memset :409), dtfeats untouched upstream -> defined 0 here only if zero_fill != 0. */
void orc_locw_feat_sum_backward(const float *lw, const float *tfeats, const float *dsum, const int32_t *lens,
                                const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn, int64_t csum,
                                double *dlw, float *dtfeats) {
    memset(dlw, 0, sizeof(double) * (size_t)(csum * m));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        for (int64_t j = 0; j < lens[i]; ++j) {
            const int64_t e = bgs[i] + j;
            for (int64_t a = 0; a < m; ++a) {
                const float w = lw[e * m + a];
                double acc = 0.0;
                for (int64_t q = 0; q < ofn; ++q) {
                    const float d = dsum[(i * m + a) * ofn + q];
                    dtfeats[(e * m + a) * ofn + q] = d * w;
                    acc += (double)d * (double)tfeats[(e * m + a) * ofn + q];
                }
                dlw[e * m + a] = acc;
            }
        }
    }
}

/* a11 — eliminateCenter TFNeighborKernelNew.cu:337-370 (writes capped at len-1, see include/pcseg_ops.h) */
void orc_eliminate_center(const int32_t *inidxs, const int32_t *lens, const int32_t *bgs, int64_t pn,
                          int32_t *onidxs, int32_t *olens, int32_t *obgs, int32_t *ocidxs) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < pn; ++i) {
        const int64_t bg = bgs[i], obg = bg - i;
        obgs[i] = (int32_t)obg;
        olens[i] = lens[i] - 1;
        int64_t w = 0;
        for (int64_t j = 0; j < lens[i]; ++j) {
            const int32_t v = inidxs[bg + j];
            if (v == (int32_t)i) continue;
            if (w < lens[i] - 1) { onidxs[obg + w] = v; ocidxs[obg + w] = (int32_t)i; }
            ++w;
        }
    }
}

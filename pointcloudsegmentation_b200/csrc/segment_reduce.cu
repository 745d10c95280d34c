// segment_reduce.cu — CSR segmented reductions: a3 NeighborSumFeatGather, a7 LocationWeightSum,
// a9 LocationWeightFeatSum, a5 NeighborMaxFeatGather, a2 NeighborGather, fp32 or bf16 storage.
//
// Two kernel families:
//  * WALKERS (this file): one group of G lanes per (centre, column chunk), channel index on the lane axis
//    (16-byte lanes when the width allows), the group walks the centre's edge rows with R independent row loads
//    in flight (ragged head of len % R rows first, then branch-free batches of R) and accumulates in registers,
//    strictly in edge order => deterministic sums, first-maximum argmax, no atomics, no memset.  R is sized so that
//    a lane keeps >= 64 bytes in flight whatever its width (R = 4 rows of 16 bytes, 8 rows of 8 / 4 bytes):
//    with 4 rows the 4-byte lanes of odd widths (C = 35, 131) had 16 bytes in flight and sat at 45-68 % of HBM.
//    93-100 % of the measured HBM peak at C = 128, 84-94 % at C = 64.
//  * STREAM (seg_stream.cu): rows narrower than a cache line.  The edge tensor of consecutive centres IS one
//    contiguous stream; a CTA brings its slice in with the TMA engine (cp.async.bulk) and reduces out of shared
//    memory.  Selected by launch_segment_* below when the row is short (pcs_set_option seg_stream).
#include "common.cuh"

using namespace pcs;

namespace pcs {
// seg_stream.cu
int launch_seg_stream_sum(bool bf, const void *feats, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                          int64_t csum, void *out, cudaStream_t s, const char *what, bool *done);
int launch_seg_stream_max(bool bf, const void *feats, const int32_t *lens, const int32_t *bgs, int64_t pn2, int64_t C,
                          int64_t pn1, void *gfeats, int32_t *max_idxs, cudaStream_t s, const char *what, bool *done);
}  // namespace pcs

namespace {
constexpr int kBlock = 256;

template <int V> __device__ __forceinline__ void sti(int32_t *p, const int (&a)[V]) {
    if constexpr (V == 1) *p = a[0];
    else if constexpr (V == 2) *reinterpret_cast<int2 *>(p) = make_int2(a[0], a[1]);
    else {
#pragma unroll
        for (int k = 0; k < V; k += 4) *reinterpret_cast<int4 *>(p + k) = make_int4(a[k], a[k + 1], a[k + 2], a[k + 3]);
    }
}

// ---- a3 / a7 / a9 segment sum (reference kernels sumFeatGather TFNeighborKernelNew.cu:4-27,
// locWSumForwardKernel TFNeighborKernel.cu:128-151, locationWeightFeatSumForwardKernel :179-211)
template <typename T, int V, int R, bool WEIGHTED>
__global__ void __launch_bounds__(kBlock)
k_segment_sum(const T *__restrict__ feats, const T *__restrict__ lw, const int32_t *__restrict__ lens,
              const int32_t *__restrict__ bgs, int64_t pn, int64_t C, int64_t m, int64_t ofn, int gshift, int CH,
              T *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const T *p = feats + bg * C + col;
    const T *wp = nullptr;
    if (WEIGHTED) wp = lw + bg * m + col / ofn;
    Vec<V> acc = vzero<V>();
    // Ragged head first (len % R rows, all loaded together), then branch-free batches of R: every step has
    // R row loads in flight and the adds stay in edge order.
    const int head = len & (R - 1);
    if (head) {
        Vec<V> v[R - 1];
        float w[R - 1];
#pragma unroll
        for (int k = 0; k < R - 1; ++k) {
            v[k] = vzero<V>(); w[k] = 0.f;
            if (k < head) { v[k] = IO<T, V>::ldg(p + k * C); if (WEIGHTED) w[k] = IO<T, 1>::ldg(wp + k * m).v[0]; }
        }
#pragma unroll
        for (int k = 0; k < R - 1; ++k)
            if (k < head) {
#pragma unroll
                for (int q = 0; q < V; ++q) acc.v[q] = WEIGHTED ? fmaf(w[k], v[k].v[q], acc.v[q]) : acc.v[q] + v[k].v[q];
            }
        p += head * C;
        if (WEIGHTED) wp += head * m;
    }
#pragma unroll 2
    for (int r = head; r < len; r += R) {
        Vec<V> v[R];
        float w[R];
#pragma unroll
        for (int k = 0; k < R; ++k) { v[k] = IO<T, V>::ldg(p + k * C); if (WEIGHTED) w[k] = IO<T, 1>::ldg(wp + k * m).v[0]; }
#pragma unroll
        for (int k = 0; k < R; ++k) {
#pragma unroll
            for (int q = 0; q < V; ++q) acc.v[q] = WEIGHTED ? fmaf(w[k], v[k].v[q], acc.v[q]) : acc.v[q] + v[k].v[q];
        }
        p += R * C;
        if (WEIGHTED) wp += R * m;
    }
    IO<T, V>::stg(out + i * C + col, acc);
}

// ---- a2 NeighborGather through the transpose CSR (reference: featGather/diffFeatGather atomics,
// TFNeighborKernel.cu:63-125).  Row n sums the edges that point AT n in ascending edge id.
template <typename T, int V, int R>
__global__ void __launch_bounds__(kBlock)
k_neighbor_gather(const T *__restrict__ sfeats, const int32_t *__restrict__ tperm, const int32_t *__restrict__ tbgs,
                  int64_t pn, int64_t C, int gshift, int CH, T *__restrict__ out) {
    int64_t n; int col;
    if (!row_item(pn, gshift, CH, V, C, n, col)) return;
    const int tb = tbgs[n], te = tbgs[n + 1];
    const T *base = sfeats + col;
    Vec<V> acc = vzero<V>();
    int p = tb;
    const int head = (te - tb) & (R - 1);
    if (head) {
        int64_t e[R - 1];
        Vec<V> v[R - 1];
#pragma unroll
        for (int k = 0; k < R - 1; ++k) e[k] = (k < head) ? (int64_t)__ldg(tperm + p + k) : 0;
#pragma unroll
        for (int k = 0; k < R - 1; ++k) { v[k] = vzero<V>(); if (k < head) v[k] = IO<T, V>::ldg(base + e[k] * C); }
#pragma unroll
        for (int k = 0; k < R - 1; ++k) if (k < head) vadd<V>(acc, v[k]);
        p += head;
    }
#pragma unroll 2
    for (; p < te; p += R) {
        int64_t e[R];
        Vec<V> v[R];
#pragma unroll
        for (int k = 0; k < R; ++k) e[k] = __ldg(tperm + p + k);
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = IO<T, V>::ldg(base + e[k] * C);
#pragma unroll
        for (int k = 0; k < R; ++k) vadd<V>(acc, v[k]);
    }
    IO<T, V>::stg(out + n * C + col, acc);
}

// use_diff: out[n] = (sum of the edges pointing at n) - (sum of n's own segment): the tperm walk, then the
// contiguous own-segment walk.  (Walking both lists in one fully predicated loop was measured 30-40 % slower,
// profiles/r3b_*: the predicates cost more than the second dependent chain.)
template <typename T, int V, int R>
__global__ void __launch_bounds__(kBlock)
k_neighbor_gather_diff(const T *__restrict__ sfeats, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
                       const int32_t *__restrict__ tperm, const int32_t *__restrict__ tbgs, int64_t pn, int64_t C,
                       int gshift, int CH, T *__restrict__ out) {
    int64_t n; int col;
    if (!row_item(pn, gshift, CH, V, C, n, col)) return;
    const int tb = tbgs[n], te = tbgs[n + 1];
    const int64_t bg = bgs[n];
    const int len = lens[n];
    const T *base = sfeats + col;
    Vec<V> acc = vzero<V>(), sub = vzero<V>();
    int p = tb;
    const int head = (te - tb) & (R - 1);
    if (head) {
        int64_t e[R - 1];
        Vec<V> v[R - 1];
#pragma unroll
        for (int k = 0; k < R - 1; ++k) e[k] = (k < head) ? (int64_t)__ldg(tperm + p + k) : 0;
#pragma unroll
        for (int k = 0; k < R - 1; ++k) { v[k] = vzero<V>(); if (k < head) v[k] = IO<T, V>::ldg(base + e[k] * C); }
#pragma unroll
        for (int k = 0; k < R - 1; ++k) if (k < head) vadd<V>(acc, v[k]);
        p += head;
    }
#pragma unroll 2
    for (; p < te; p += R) {
        int64_t e[R];
        Vec<V> v[R];
#pragma unroll
        for (int k = 0; k < R; ++k) e[k] = __ldg(tperm + p + k);
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = IO<T, V>::ldg(base + e[k] * C);
#pragma unroll
        for (int k = 0; k < R; ++k) vadd<V>(acc, v[k]);
    }
    const T *q = base + bg * C;
    const int h2 = len & (R - 1);
    if (h2) {
        Vec<V> v[R - 1];
#pragma unroll
        for (int k = 0; k < R - 1; ++k) { v[k] = vzero<V>(); if (k < h2) v[k] = IO<T, V>::ldg(q + k * C); }
#pragma unroll
        for (int k = 0; k < R - 1; ++k) if (k < h2) vadd<V>(sub, v[k]);
        q += h2 * C;
    }
#pragma unroll 2
    for (int r = h2; r < len; r += R) {
        Vec<V> v[R];
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = IO<T, V>::ldg(q + k * C);
#pragma unroll
        for (int k = 0; k < R; ++k) vadd<V>(sub, v[k]);
        q += R * C;
    }
    IO<T, V>::stg(out + n * C + col, vsub<V>(acc, sub));
}

// a5: per-segment channel max + offset of the first maximum; strict '<' exactly as
// maxFeatGather TFNeighborKernelNew.cu:83-118 (a NaN is selected only if it is the first row).
template <typename T, int V, int R>
__global__ void __launch_bounds__(kBlock)
k_max_gather(const T *__restrict__ feats, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
             int64_t pn2, int64_t C, int gshift, int CH, T *__restrict__ gfeats, int32_t *__restrict__ max_idxs) {
    int64_t i; int col;
    if (!row_item(pn2, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<V> best = vzero<V>();
    int bi[V];
#pragma unroll
    for (int k = 0; k < V; ++k) bi[k] = 0;
    if (len > 0) {
        const T *p = feats + bg * C + col;
        // ragged head of 1..R rows (row 0 seeds the max), then branch-free batches of R
        const int head = ((len - 1) & (R - 1)) + 1;
        {
            Vec<V> v[R];
#pragma unroll
            for (int k = 0; k < R; ++k) if (k < head) v[k] = IO<T, V>::ldg(p + k * C);
            best = v[0];
#pragma unroll
            for (int k = 1; k < R; ++k)
                if (k < head) {
#pragma unroll
                    for (int q = 0; q < V; ++q)
                        if (best.v[q] < v[k].v[q]) { best.v[q] = v[k].v[q]; bi[q] = k; }
                }
            p += head * C;
        }
#pragma unroll 2
        for (int r = head; r < len; r += R) {
            Vec<V> v[R];
#pragma unroll
            for (int k = 0; k < R; ++k) v[k] = IO<T, V>::ldg(p + k * C);
#pragma unroll
            for (int k = 0; k < R; ++k) {
#pragma unroll
                for (int q = 0; q < V; ++q)
                    if (best.v[q] < v[k].v[q]) { best.v[q] = v[k].v[q]; bi[q] = r + k; }
            }
            p += R * C;
        }
    }
    IO<T, V>::stg(gfeats + i * C + col, best);
    sti<V>(max_idxs + i * C + col, bi);
}

// rows in flight per lane: 64 bytes' worth, at least 4, at most 8 (pcs_set_option walk_rows_narrow overrides
// the narrow-lane value with 4 / 8 / 16 for A/B runs)
template <typename T, int V> constexpr int auto_rows() { return (V * (int)sizeof(T) >= 16) ? 4 : 8; }
}  // namespace

#define PCS_ROWS_SWITCH(T_, V_, ...)                                                                    \
    do {                                                                                                \
        const int64_t rr_ = (V_ * (int)sizeof(T_) >= 16) ? 4 : opt::walk_rows_narrow.load();             \
        if (rr_ >= 16) { constexpr int R = 16; __VA_ARGS__; }                                            \
        else if (rr_ >= 8) { constexpr int R = 8; __VA_ARGS__; }                                         \
        else { constexpr int R = 4; __VA_ARGS__; }                                                       \
    } while (0)

static int launch_segment_sum(bool bf, const void *feats, const void *lw, const int32_t *lens, const int32_t *bgs,
                              int64_t pn, int64_t C, int64_t m, int64_t ofn, int64_t csum, void *out, cudaStream_t s,
                              const char *what) {
    const bool weighted = lw != nullptr;
    const int es = bf ? 2 : 4;
    if (!weighted) {  // short rows: TMA-staged stream reduction
        bool done = false;
        if (launch_seg_stream_sum(bf, feats, lens, bgs, pn, C, csum, out, s, what, &done)) return 1;
        if (done) return 0;
    }
    int align = common_align({feats, out});
    RowTiling t = make_tiling_es(C, es, align);
    if (weighted) {  // a lane's V columns must share one anchor: V | ofn
        while (t.V > 1 && ofn % t.V) t.V >>= 1;
        t = make_tiling_es(C, es, std::min(align, t.V * es));
    }
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, PCS_ROWS_SWITCH(T, V,
        if (weighted) k_segment_sum<T, V, (R > 8 ? 8 : R), true><<<grid, kBlock, 0, s>>>((const T *)feats, (const T *)lw, lens, bgs, pn, C, m, ofn, t.gshift, t.CH, (T *)out);
        else k_segment_sum<T, V, R, false><<<grid, kBlock, 0, s>>>((const T *)feats, (const T *)lw, lens, bgs, pn, C, m, ofn, t.gshift, t.CH, (T *)out)));
    return check_launch(what);
}

static int neighbor_gather_impl(bool bf, const void *sfeats, const int32_t *lens, const int32_t *bgs, const int32_t *tperm,
                                const int32_t *tbgs, int64_t pn, int64_t ifn, int64_t csum, int use_diff, void *ifeats,
                                pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_gather: negative size");
    if (pn == 0 || ifn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    RowTiling t = make_tiling_es(ifn, bf ? 2 : 4, common_align({sfeats, ifeats}));
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, PCS_ROWS_SWITCH(T, V,
        if (use_diff) k_neighbor_gather_diff<T, V, (R > 8 ? 8 : R)><<<grid, kBlock, 0, s>>>((const T *)sfeats, lens, bgs, tperm, tbgs, pn, ifn, t.gshift, t.CH, (T *)ifeats);
        else k_neighbor_gather<T, V, R><<<grid, kBlock, 0, s>>>((const T *)sfeats, tperm, tbgs, pn, ifn, t.gshift, t.CH, (T *)ifeats)));
    return check_launch("neighbor_gather");
}

static int max_feat_gather_impl(bool bf, const void *ifeats, const int32_t *vlens, const int32_t *vbegs, int64_t pn1,
                                int64_t pn2, int64_t fd, void *gfeats, int32_t *max_idxs, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0 && fd >= 0, "neighbor_max_feat_gather: negative size");
    if (pn2 == 0 || fd == 0) return 0;
    cudaStream_t s = as_stream(stream);
    bool done = false;
    if (launch_seg_stream_max(bf, ifeats, vlens, vbegs, pn2, fd, pn1, gfeats, max_idxs, s, "neighbor_max_feat_gather", &done)) return 1;
    if (done) return 0;
    const int es = bf ? 2 : 4;
    int align = common_align({ifeats, gfeats});
    // the int32 arg-max row is written with the same lane width: V ints must stay 16-byte-chunk aligned
    int ia = common_align({max_idxs});
    align = std::min(align, std::max(es, ia * es / 4));
    RowTiling t = make_tiling_es(fd, es, align);
    int64_t grid = grid_for(pn2, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, PCS_ROWS_SWITCH(T, V,
        k_max_gather<T, V, (R > 8 ? 8 : R)><<<grid, kBlock, 0, s>>>((const T *)ifeats, vlens, vbegs, pn2, fd, t.gshift, t.CH, (T *)gfeats, max_idxs)));
    return check_launch("neighbor_max_feat_gather");
}

// ---- C ABI: fp32 entries (one per reference launcher) and their bf16-storage twins ------------------------
extern "C" int pcs_neighbor_sum_feat_gather(const float *ifeats, const int32_t *lens, const int32_t *bgs,
                                            int64_t pn, int64_t fd, int64_t csum, float *ogfeats,
                                            pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fd >= 0 && csum >= 0, "neighbor_sum_feat_gather: negative size");
    if (pn == 0 || fd == 0) return 0;
    return launch_segment_sum(false, ifeats, nullptr, lens, bgs, pn, fd, 0, 1, csum, ogfeats, as_stream(stream),
                              "neighbor_sum_feat_gather");
}
extern "C" int pcs_neighbor_sum_feat_gather_bf16(const void *ifeats, const int32_t *lens, const int32_t *bgs,
                                                 int64_t pn, int64_t fd, int64_t csum, void *ogfeats,
                                                 pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fd >= 0 && csum >= 0, "neighbor_sum_feat_gather: negative size");
    if (pn == 0 || fd == 0) return 0;
    return launch_segment_sum(true, ifeats, nullptr, lens, bgs, pn, fd, 0, 1, csum, ogfeats, as_stream(stream),
                              "neighbor_sum_feat_gather");
}

extern "C" int pcs_location_weight_sum(const float *lw, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                       int64_t m, int64_t csum, float *lw_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && csum >= 0, "location_weight_sum: negative size");
    if (pn == 0 || m == 0) return 0;
    return launch_segment_sum(false, lw, nullptr, lens, bgs, pn, m, 0, 1, csum, lw_sum, as_stream(stream),
                              "location_weight_sum");
}
extern "C" int pcs_location_weight_sum_bf16(const void *lw, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                            int64_t m, int64_t csum, void *lw_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && csum >= 0, "location_weight_sum: negative size");
    if (pn == 0 || m == 0) return 0;
    return launch_segment_sum(true, lw, nullptr, lens, bgs, pn, m, 0, 1, csum, lw_sum, as_stream(stream),
                              "location_weight_sum");
}

extern "C" int pcs_location_weight_feat_sum(const float *lw, const float *tfeats, const int32_t *lens,
                                            const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn,
                                            int64_t csum, float *otfeats_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum: negative size");
    if (pn == 0 || m == 0 || ofn == 0) return 0;
    return launch_segment_sum(false, tfeats, lw, lens, bgs, pn, m * ofn, m, ofn, csum, otfeats_sum, as_stream(stream),
                              "location_weight_feat_sum");
}
extern "C" int pcs_location_weight_feat_sum_bf16(const void *lw, const void *tfeats, const int32_t *lens,
                                                 const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn,
                                                 int64_t csum, void *otfeats_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum: negative size");
    if (pn == 0 || m == 0 || ofn == 0) return 0;
    return launch_segment_sum(true, tfeats, lw, lens, bgs, pn, m * ofn, m, ofn, csum, otfeats_sum, as_stream(stream),
                              "location_weight_feat_sum");
}

extern "C" int pcs_neighbor_gather(const float *sfeats, const int32_t *lens, const int32_t *bgs,
                                   const int32_t *tperm, const int32_t *tbgs, int64_t pn, int64_t ifn,
                                   int64_t csum, int use_diff, float *ifeats, pcs_stream_t stream) {
    return neighbor_gather_impl(false, sfeats, lens, bgs, tperm, tbgs, pn, ifn, csum, use_diff, ifeats, stream);
}
extern "C" int pcs_neighbor_gather_bf16(const void *sfeats, const int32_t *lens, const int32_t *bgs,
                                        const int32_t *tperm, const int32_t *tbgs, int64_t pn, int64_t ifn,
                                        int64_t csum, int use_diff, void *ifeats, pcs_stream_t stream) {
    return neighbor_gather_impl(true, sfeats, lens, bgs, tperm, tbgs, pn, ifn, csum, use_diff, ifeats, stream);
}

extern "C" int pcs_neighbor_max_feat_gather(const float *ifeats, const int32_t *vlens, const int32_t *vbegs,
                                            int64_t pn1, int64_t pn2, int64_t fd, float *gfeats,
                                            int32_t *max_idxs, pcs_stream_t stream) {
    return max_feat_gather_impl(false, ifeats, vlens, vbegs, pn1, pn2, fd, gfeats, max_idxs, stream);
}
extern "C" int pcs_neighbor_max_feat_gather_bf16(const void *ifeats, const int32_t *vlens, const int32_t *vbegs,
                                                 int64_t pn1, int64_t pn2, int64_t fd, void *gfeats,
                                                 int32_t *max_idxs, pcs_stream_t stream) {
    return max_feat_gather_impl(true, ifeats, vlens, vbegs, pn1, pn2, fd, gfeats, max_idxs, stream);
}

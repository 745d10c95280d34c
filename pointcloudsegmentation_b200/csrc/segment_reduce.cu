// segment_reduce.cu — CSR segmented reductions: a3 NeighborSumFeatGather, a7 LocationWeightSum,
// a9 LocationWeightFeatSum, a5 NeighborMaxFeatGather, a2 NeighborGather.
// One group of G lanes per (centre, column chunk): channel index on the lane axis (16-byte lanes when
// C % 4 == 0), the group walks the centre's edge rows with 4-8 independent row loads in flight (ragged
// head of len % 4 rows first, then branch-free batches of 4) and accumulates in registers, strictly in
// edge order => deterministic sums, first-maximum argmax, no atomics, no memset.  Measured against a ring-buffered "streaming" variant that prefetched across
// segment boundaries (profiles/r1d_bench_ops_{walk,stream}_kernels.log): the walking kernels reach
// 93-100 % of the measured HBM peak at C=128 and 78-90 % at C=64, the streaming variant 35-62 % — its
// per-row cursor bookkeeping made it issue-bound — so the walking form is the one kept.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// ---- a3 / a7 / a9 segment sum (reference kernels sumFeatGather TFNeighborKernelNew.cu:4-27,
// locWSumForwardKernel TFNeighborKernel.cu:128-151, locationWeightFeatSumForwardKernel :179-211)
template <int V, bool WEIGHTED>
__global__ void __launch_bounds__(kBlock)
k_segment_sum(const float *__restrict__ feats, const float *__restrict__ lw,
              const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn, int64_t C,
              int64_t m, int64_t ofn, int gshift, int CH, float *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const float *p = feats + bg * C + col;
    const float *wp = nullptr;
    if (WEIGHTED) wp = lw + bg * m + col / ofn;
    Vec<V> acc = vzero<V>();
    // Ragged head first (len % 4 rows, all loaded together), then branch-free batches of 4: every step
    // has several row loads in flight and the adds stay in edge order.  (Walking the remainder one row
    // = one memory latency at a time cost 1-3 extra round trips per centre; predicating the batch
    // itself made the main loop slower, profiles/README.md r1l.)
    const int head = len & 3;
    if (head) {
        Vec<V> v0 = ld<V>(p), v1 = vzero<V>(), v2 = vzero<V>();
        if (head > 1) v1 = ld<V>(p + C);
        if (head > 2) v2 = ld<V>(p + 2 * C);
        if (WEIGHTED) {
            float w0 = __ldg(wp), w1 = 0.f, w2 = 0.f;
            if (head > 1) w1 = __ldg(wp + m);
            if (head > 2) w2 = __ldg(wp + 2 * m);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                acc.v[k] = fmaf(w0, v0.v[k], acc.v[k]);
                if (head > 1) acc.v[k] = fmaf(w1, v1.v[k], acc.v[k]);
                if (head > 2) acc.v[k] = fmaf(w2, v2.v[k], acc.v[k]);
            }
            wp += head * m;
        } else {
            vadd<V>(acc, v0);
            if (head > 1) vadd<V>(acc, v1);
            if (head > 2) vadd<V>(acc, v2);
        }
        p += head * C;
    }
#pragma unroll 2
    for (int r = head; r < len; r += 4) {
        Vec<V> v0 = ld<V>(p), v1 = ld<V>(p + C), v2 = ld<V>(p + 2 * C), v3 = ld<V>(p + 3 * C);
        if (WEIGHTED) {
            float w0 = __ldg(wp), w1 = __ldg(wp + m), w2 = __ldg(wp + 2 * m), w3 = __ldg(wp + 3 * m);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                acc.v[k] = fmaf(w0, v0.v[k], acc.v[k]);
                acc.v[k] = fmaf(w1, v1.v[k], acc.v[k]);
                acc.v[k] = fmaf(w2, v2.v[k], acc.v[k]);
                acc.v[k] = fmaf(w3, v3.v[k], acc.v[k]);
            }
            wp += 4 * m;
        } else {
            vadd<V>(acc, v0); vadd<V>(acc, v1); vadd<V>(acc, v2); vadd<V>(acc, v3);
        }
        p += 4 * C;
    }
    st<V>(out + i * C + col, acc);
}

// ---- a2 NeighborGather through the transpose CSR (reference: featGather/diffFeatGather atomics,
// TFNeighborKernel.cu:63-125).  Row n sums the edges that point AT n in ascending edge id.
template <int V, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_neighbor_gather(const float *__restrict__ sfeats, const int32_t *__restrict__ lens,
                  const int32_t *__restrict__ bgs, const int32_t *__restrict__ tperm,
                  const int32_t *__restrict__ tbgs, int64_t pn, int64_t C, int gshift, int CH,
                  float *__restrict__ out) {
    int64_t n; int col;
    if (!row_item(pn, gshift, CH, V, C, n, col)) return;
    const int tb = tbgs[n], te = tbgs[n + 1];
    const float *base = sfeats + col;
    Vec<V> acc = vzero<V>();
    int p = tb;
    const int head = (te - tb) & 3;
    if (head) {  // ragged head: up to 3 gathers issued together
        int64_t e0 = tperm[p], e1 = e0, e2 = e0;
        if (head > 1) e1 = tperm[p + 1];
        if (head > 2) e2 = tperm[p + 2];
        Vec<V> v0 = ld<V>(base + e0 * C), v1 = vzero<V>(), v2 = vzero<V>();
        if (head > 1) v1 = ld<V>(base + e1 * C);
        if (head > 2) v2 = ld<V>(base + e2 * C);
        vadd<V>(acc, v0);
        if (head > 1) vadd<V>(acc, v1);
        if (head > 2) vadd<V>(acc, v2);
        p += head;
    }
#pragma unroll 2
    for (; p < te; p += 4) {
        int64_t e0 = tperm[p], e1 = tperm[p + 1], e2 = tperm[p + 2], e3 = tperm[p + 3];
        Vec<V> v0 = ld<V>(base + e0 * C), v1 = ld<V>(base + e1 * C);
        Vec<V> v2 = ld<V>(base + e2 * C), v3 = ld<V>(base + e3 * C);
        vadd<V>(acc, v0); vadd<V>(acc, v1); vadd<V>(acc, v2); vadd<V>(acc, v3);
    }
    if (DIFF) {
        const int64_t bg = bgs[n];
        const int len = lens[n];
        const float *q = base + bg * C;
        Vec<V> sub = vzero<V>();
        const int h2 = len & 3;
        if (h2) {
            Vec<V> v0 = ld<V>(q), v1 = vzero<V>(), v2 = vzero<V>();
            if (h2 > 1) v1 = ld<V>(q + C);
            if (h2 > 2) v2 = ld<V>(q + 2 * C);
            vadd<V>(sub, v0);
            if (h2 > 1) vadd<V>(sub, v1);
            if (h2 > 2) vadd<V>(sub, v2);
            q += h2 * C;
        }
#pragma unroll 2
        for (int r = h2; r < len; r += 4) {
            Vec<V> v0 = ld<V>(q), v1 = ld<V>(q + C), v2 = ld<V>(q + 2 * C), v3 = ld<V>(q + 3 * C);
            vadd<V>(sub, v0); vadd<V>(sub, v1); vadd<V>(sub, v2); vadd<V>(sub, v3);
            q += 4 * C;
        }
        acc = vsub<V>(acc, sub);
    }
    st<V>(out + n * C + col, acc);
}
// a5: per-segment channel max + offset of the first maximum; strict '<' exactly as
// maxFeatGather TFNeighborKernelNew.cu:83-118 (a NaN is selected only if it is the first row).
template <int V>
__global__ void __launch_bounds__(kBlock)
k_max_gather(const float *__restrict__ feats, const int32_t *__restrict__ lens,
             const int32_t *__restrict__ bgs, int64_t pn2, int64_t C, int gshift, int CH,
             float *__restrict__ gfeats, int32_t *__restrict__ max_idxs) {
    int64_t i; int col;
    if (!row_item(pn2, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<V> best = vzero<V>();
    int bi[V];
#pragma unroll
    for (int k = 0; k < V; ++k) bi[k] = 0;
    if (len > 0) {
        const float *p = feats + bg * C + col;
        // ragged head of 1..4 rows (row 0 seeds the max), then branch-free batches of 4
        const int head = ((len - 1) & 3) + 1;
        {
            Vec<V> v0 = ld<V>(p), v1 = v0, v2 = v0, v3 = v0;
            if (head > 1) v1 = ld<V>(p + C);
            if (head > 2) v2 = ld<V>(p + 2 * C);
            if (head > 3) v3 = ld<V>(p + 3 * C);
            best = v0;
#pragma unroll
            for (int k = 0; k < V; ++k) {
                if (head > 1 && best.v[k] < v1.v[k]) { best.v[k] = v1.v[k]; bi[k] = 1; }
                if (head > 2 && best.v[k] < v2.v[k]) { best.v[k] = v2.v[k]; bi[k] = 2; }
                if (head > 3 && best.v[k] < v3.v[k]) { best.v[k] = v3.v[k]; bi[k] = 3; }
            }
            p += head * C;
        }
#pragma unroll 2
        for (int r = head; r < len; r += 4) {
            Vec<V> v0 = ld<V>(p), v1 = ld<V>(p + C), v2 = ld<V>(p + 2 * C), v3 = ld<V>(p + 3 * C);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                if (best.v[k] < v0.v[k]) { best.v[k] = v0.v[k]; bi[k] = r; }
                if (best.v[k] < v1.v[k]) { best.v[k] = v1.v[k]; bi[k] = r + 1; }
                if (best.v[k] < v2.v[k]) { best.v[k] = v2.v[k]; bi[k] = r + 2; }
                if (best.v[k] < v3.v[k]) { best.v[k] = v3.v[k]; bi[k] = r + 3; }
            }
            p += 4 * C;
        }
    }
    st<V>(gfeats + i * C + col, best);
    Vec<V> bv;
#pragma unroll
    for (int k = 0; k < V; ++k) bv.v[k] = __int_as_float(bi[k]);
    st<V>(reinterpret_cast<float *>(max_idxs) + i * C + col, bv);
}

}  // namespace

#define PCS_DISPATCH_V(V_, ...)          \
    switch (V_) {                         \
        case 4: { constexpr int V = 4; __VA_ARGS__; } break; \
        case 2: { constexpr int V = 2; __VA_ARGS__; } break; \
        default: { constexpr int V = 1; __VA_ARGS__; } break; \
    }

static int launch_segment_sum(const float *feats, const float *lw, const int32_t *lens, const int32_t *bgs,
                              int64_t pn, int64_t C, int64_t m, int64_t ofn, float *out, cudaStream_t s,
                              const char *what) {
    bool weighted = lw != nullptr;
    bool can4 = aligned(feats, 16) && aligned(out, 16) && (!weighted || ofn % 4 == 0);
    bool can2 = aligned(feats, 8) && aligned(out, 8) && (!weighted || ofn % 2 == 0);
    RowTiling t = make_tiling(C, can4, can2);
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_V(t.V,
        if (weighted) k_segment_sum<V, true><<<grid, kBlock, 0, s>>>(feats, lw, lens, bgs, pn, C, m, ofn, t.gshift, t.CH, out);
        else k_segment_sum<V, false><<<grid, kBlock, 0, s>>>(feats, lw, lens, bgs, pn, C, m, ofn, t.gshift, t.CH, out));
    return check_launch(what);
}

extern "C" int pcs_neighbor_sum_feat_gather(const float *ifeats, const int32_t *lens, const int32_t *bgs,
                                            int64_t pn, int64_t fd, int64_t csum, float *ogfeats,
                                            pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fd >= 0 && csum >= 0, "neighbor_sum_feat_gather: negative size");
    if (pn == 0 || fd == 0) return 0;
    return launch_segment_sum(ifeats, nullptr, lens, bgs, pn, fd, 0, 1, ogfeats, as_stream(stream),
                              "neighbor_sum_feat_gather");
}

extern "C" int pcs_location_weight_sum(const float *lw, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                       int64_t m, int64_t csum, float *lw_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && csum >= 0, "location_weight_sum: negative size");
    if (pn == 0 || m == 0) return 0;
    return launch_segment_sum(lw, nullptr, lens, bgs, pn, m, 0, 1, lw_sum, as_stream(stream),
                              "location_weight_sum");
}

extern "C" int pcs_location_weight_feat_sum(const float *lw, const float *tfeats, const int32_t *lens,
                                            const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn,
                                            int64_t csum, float *otfeats_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum: negative size");
    if (pn == 0 || m == 0 || ofn == 0) return 0;
    return launch_segment_sum(tfeats, lw, lens, bgs, pn, m * ofn, m, ofn, otfeats_sum, as_stream(stream),
                              "location_weight_feat_sum");
}

extern "C" int pcs_neighbor_gather(const float *sfeats, const int32_t *lens, const int32_t *bgs,
                                   const int32_t *tperm, const int32_t *tbgs, int64_t pn, int64_t ifn,
                                   int64_t csum, int use_diff, float *ifeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_gather: negative size");
    if (pn == 0 || ifn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    RowTiling t = make_tiling(ifn, aligned(sfeats, 16) && aligned(ifeats, 16), aligned(sfeats, 8) && aligned(ifeats, 8));
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_V(t.V,
        if (use_diff) k_neighbor_gather<V, true><<<grid, kBlock, 0, s>>>(sfeats, lens, bgs, tperm, tbgs, pn, ifn, t.gshift, t.CH, ifeats);
        else k_neighbor_gather<V, false><<<grid, kBlock, 0, s>>>(sfeats, lens, bgs, tperm, tbgs, pn, ifn, t.gshift, t.CH, ifeats));
    return check_launch("neighbor_gather");
}

extern "C" int pcs_neighbor_max_feat_gather(const float *ifeats, const int32_t *vlens, const int32_t *vbegs,
                                            int64_t pn1, int64_t pn2, int64_t fd, float *gfeats,
                                            int32_t *max_idxs, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0 && fd >= 0, "neighbor_max_feat_gather: negative size");
    if (pn2 == 0 || fd == 0) return 0;
    bool a16 = aligned(ifeats, 16) && aligned(gfeats, 16) && aligned(max_idxs, 16);
    bool a8 = aligned(ifeats, 8) && aligned(gfeats, 8) && aligned(max_idxs, 8);
    RowTiling t = make_tiling(fd, a16, a8);
    int64_t grid = grid_for(pn2, t, kBlock);
    cudaStream_t s = as_stream(stream);
    PCS_DISPATCH_V(t.V, k_max_gather<V><<<grid, kBlock, 0, s>>>(ifeats, vlens, vbegs, pn2, fd, t.gshift, t.CH, gfeats, max_idxs));
    return check_launch("neighbor_max_feat_gather");
}

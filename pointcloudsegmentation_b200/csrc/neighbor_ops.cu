// neighbor_ops.cu — CSR neighbour operators a1..a8 for sm_100a.
// Design: channel index on the lane axis (coalesced 128 B rows), a group of G lanes owns one
// (centre, column-chunk) work item and walks the centre's edges with 4 independent row loads in
// flight; every reduction is a sequential per-lane accumulation in edge order => deterministic,
// no atomics (the reference's featGather/diffFeatGather use fp32 atomicAdd,
// TFNeighborKernel.cu:83,112,121).
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// ---- a1 NeighborScatter (reference kernels featScatter/diffFeatScatter, TFNeighborKernel.cu:5-59)
template <int V, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_neighbor_scatter(const float *__restrict__ ifeats, const int32_t *__restrict__ nidxs,
                   const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn,
                   int64_t C, int gshift, int CH, float *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<V> c = vzero<V>();
    if (DIFF) c = ld<V>(ifeats + i * C + col);
    const int32_t *np = nidxs + bg;
    float *op = out + bg * C + col;
    int r = 0;
    for (; r + 4 <= len; r += 4) {
        int64_t n0 = np[r], n1 = np[r + 1], n2 = np[r + 2], n3 = np[r + 3];
        Vec<V> v0 = ld<V>(ifeats + n0 * C + col);
        Vec<V> v1 = ld<V>(ifeats + n1 * C + col);
        Vec<V> v2 = ld<V>(ifeats + n2 * C + col);
        Vec<V> v3 = ld<V>(ifeats + n3 * C + col);
        st<V>(op, vsub<V>(v0, c));
        st<V>(op + C, vsub<V>(v1, c));
        st<V>(op + 2 * C, vsub<V>(v2, c));
        st<V>(op + 3 * C, vsub<V>(v3, c));
        op += 4 * C;
    }
    for (; r < len; ++r) {
        int64_t n0 = np[r];
        st<V>(op, vsub<V>(ld<V>(ifeats + n0 * C + col), c));
        op += C;
    }
}

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
    int r = 0;
    for (; r + 4 <= len; r += 4) {
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
    for (; r < len; ++r) {
        Vec<V> v0 = ld<V>(p);
        if (WEIGHTED) {
            float w0 = __ldg(wp);
#pragma unroll
            for (int k = 0; k < V; ++k) acc.v[k] = fmaf(w0, v0.v[k], acc.v[k]);
            wp += m;
        } else {
            vadd<V>(acc, v0);
        }
        p += C;
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
    for (; p + 4 <= te; p += 4) {
        int64_t e0 = tperm[p], e1 = tperm[p + 1], e2 = tperm[p + 2], e3 = tperm[p + 3];
        Vec<V> v0 = ld<V>(base + e0 * C), v1 = ld<V>(base + e1 * C);
        Vec<V> v2 = ld<V>(base + e2 * C), v3 = ld<V>(base + e3 * C);
        vadd<V>(acc, v0); vadd<V>(acc, v1); vadd<V>(acc, v2); vadd<V>(acc, v3);
    }
    for (; p < te; ++p) vadd<V>(acc, ld<V>(base + (int64_t)tperm[p] * C));
    if (DIFF) {
        const int64_t bg = bgs[n];
        const int len = lens[n];
        const float *q = base + bg * C;
        Vec<V> sub = vzero<V>();
        int r = 0;
        for (; r + 4 <= len; r += 4) {
            Vec<V> v0 = ld<V>(q), v1 = ld<V>(q + C), v2 = ld<V>(q + 2 * C), v3 = ld<V>(q + 3 * C);
            vadd<V>(sub, v0); vadd<V>(sub, v1); vadd<V>(sub, v2); vadd<V>(sub, v3);
            q += 4 * C;
        }
        for (; r < len; ++r) { vadd<V>(sub, ld<V>(q)); q += C; }
        acc = vsub<V>(acc, sub);
    }
    st<V>(out + n * C + col, acc);
}
}  // namespace

// ---- host side ---------------------------------------------------------------------------------
#define PCS_DISPATCH_V(V_, ...)          \
    switch (V_) {                         \
        case 4: { constexpr int V = 4; __VA_ARGS__; } break; \
        case 2: { constexpr int V = 2; __VA_ARGS__; } break; \
        default: { constexpr int V = 1; __VA_ARGS__; } break; \
    }

extern "C" int pcs_neighbor_scatter(const float *ifeats, const int32_t *nidxs, const int32_t *lens,
                                    const int32_t *bgs, int64_t pn, int64_t ifn, int64_t csum,
                                    int use_diff, float *sfeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_scatter: negative size");
    if (pn == 0 || ifn == 0 || csum == 0) return 0;
    RowTiling t = make_tiling(ifn, aligned(ifeats, 16) && aligned(sfeats, 16),
                              aligned(ifeats, 8) && aligned(sfeats, 8));
    int64_t grid = grid_for(pn, t, kBlock);
    cudaStream_t s = as_stream(stream);
    PCS_DISPATCH_V(t.V,
        if (use_diff) k_neighbor_scatter<V, true><<<grid, kBlock, 0, s>>>(ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, sfeats);
        else k_neighbor_scatter<V, false><<<grid, kBlock, 0, s>>>(ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, sfeats));
    return check_launch("neighbor_scatter");
}

static int launch_segment_sum(const float *feats, const float *lw, const int32_t *lens,
                              const int32_t *bgs, int64_t pn, int64_t C, int64_t m, int64_t ofn,
                              float *out, cudaStream_t s, const char *what) {
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
    RowTiling t = make_tiling(ifn, aligned(sfeats, 16) && aligned(ifeats, 16),
                              aligned(sfeats, 8) && aligned(ifeats, 8));
    int64_t grid = grid_for(pn, t, kBlock);
    cudaStream_t s = as_stream(stream);
    PCS_DISPATCH_V(t.V,
        if (use_diff) k_neighbor_gather<V, true><<<grid, kBlock, 0, s>>>(sfeats, lens, bgs, tperm, tbgs, pn, ifn, t.gshift, t.CH, ifeats);
        else k_neighbor_gather<V, false><<<grid, kBlock, 0, s>>>(sfeats, lens, bgs, tperm, tbgs, pn, ifn, t.gshift, t.CH, ifeats));
    return check_launch("neighbor_gather");
}

// neighbor_ops.cu — a1 NeighborScatter entry + its centre-walking fallback kernel.
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
    const int head = len & 3;
    if (head) {  // ragged head: up to 3 gathers issued together, then branch-free batches of 4
        int64_t n0 = np[0], n1 = n0, n2 = n0;
        if (head > 1) n1 = np[1];
        if (head > 2) n2 = np[2];
        Vec<V> v0 = ld<V>(ifeats + n0 * C + col), v1 = v0, v2 = v0;
        if (head > 1) v1 = ld<V>(ifeats + n1 * C + col);
        if (head > 2) v2 = ld<V>(ifeats + n2 * C + col);
        st<V>(op, vsub<V>(v0, c));
        if (head > 1) st<V>(op + C, vsub<V>(v1, c));
        if (head > 2) st<V>(op + 2 * C, vsub<V>(v2, c));
        op += head * C;
    }
#pragma unroll 2
    for (int r = head; r < len; r += 4) {
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
}

}  // namespace

// ---- host side ---------------------------------------------------------------------------------
#define PCS_DISPATCH_V(V_, ...)          \
    switch (V_) {                         \
        case 4: { constexpr int V = 4; __VA_ARGS__; } break; \
        case 2: { constexpr int V = 2; __VA_ARGS__; } break; \
        default: { constexpr int V = 1; __VA_ARGS__; } break; \
    }

namespace pcs {
int launch_flat_rows(const float *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                     int64_t C, float *out, cudaStream_t s, const char *what);
}

extern "C" int pcs_neighbor_scatter(const float *ifeats, const int32_t *nidxs, const int32_t *lens,
                                    const int32_t *bgs, const int32_t *cidxs, int64_t pn, int64_t ifn,
                                    int64_t csum, int use_diff, float *sfeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_scatter: negative size");
    if (pn == 0 || ifn == 0 || csum == 0) return 0;
    cudaStream_t s = as_stream(stream);
    // Two kernels: the edge-parallel flat one (needs the centre of every edge, cidxs, for use_diff) and the
    // centre-walking one (centre row in registers, one index->gather chain per 4 rows).  The walk wins for
    // wide rows; thresholds are tunable (pcs_set_option) and their defaults come from tools/ab_rows.py.
    const bool vec4 = (ifn % 4 == 0) && aligned(ifeats, 16);
    const int64_t walk_min = use_diff ? (vec4 ? opt::scatter_walk_min_c_diff.load() : opt::scatter_walk_min_c_diff_odd.load())
                                      : (vec4 ? opt::scatter_walk_min_c_copy.load() : opt::scatter_walk_min_c_copy_odd.load());
    const bool walk = ifn >= walk_min;
    if (aligned(sfeats, 16) && !walk && (!use_diff || cidxs))
        return launch_flat_rows(ifeats, pn, nidxs, use_diff ? cidxs : nullptr, csum, ifn, sfeats, s, "neighbor_scatter");
    RowTiling t = make_tiling(ifn, aligned(ifeats, 16) && aligned(sfeats, 16),
                              aligned(ifeats, 8) && aligned(sfeats, 8));
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_V(t.V,
        if (use_diff) k_neighbor_scatter<V, true><<<grid, kBlock, 0, s>>>(ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, sfeats);
        else k_neighbor_scatter<V, false><<<grid, kBlock, 0, s>>>(ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, sfeats));
    return check_launch("neighbor_scatter");
}

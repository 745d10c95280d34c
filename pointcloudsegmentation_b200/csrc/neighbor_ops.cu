// neighbor_ops.cu — a1 NeighborScatter entry + its centre-walking kernel; fp32 or bf16 storage.
// Design: channel index on the lane axis (coalesced rows), a group of G lanes owns one (centre, column-chunk)
// work item and walks the centre's edges with R independent index->gather chains in flight (R = 4 rows of
// 16-byte lanes, 8 rows of narrower ones: >= 64 bytes in flight per lane); the centre row stays in registers.
// The edge-parallel alternative for narrow rows is flat_rows.cu.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// ---- a1 NeighborScatter (reference kernels featScatter/diffFeatScatter, TFNeighborKernel.cu:5-59)
template <typename T, int V, int R, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_neighbor_scatter(const T *__restrict__ ifeats, const int32_t *__restrict__ nidxs,
                   const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn,
                   int64_t C, int gshift, int CH, T *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<V> c = vzero<V>();
    if (DIFF) c = IO<T, V>::ldg(ifeats + i * C + col);
    const int32_t *np = nidxs + bg;
    T *op = out + bg * C + col;
    const T *src = ifeats + col;
    const int head = len & (R - 1);
    if (head) {  // ragged head: up to R-1 gathers issued together, then branch-free batches of R
        int64_t n[R - 1];
        Vec<V> v[R - 1];
#pragma unroll
        for (int k = 0; k < R - 1; ++k) n[k] = (k < head) ? (int64_t)__ldg(np + k) : 0;
#pragma unroll
        for (int k = 0; k < R - 1; ++k) if (k < head) v[k] = IO<T, V>::ldg(src + n[k] * C);
#pragma unroll
        for (int k = 0; k < R - 1; ++k) if (k < head) IO<T, V>::stg(op + k * C, DIFF ? vsub<V>(v[k], c) : v[k]);
        op += head * C;
        np += head;
    }
#pragma unroll 2
    for (int r = head; r < len; r += R) {
        int64_t n[R];
        Vec<V> v[R];
#pragma unroll
        for (int k = 0; k < R; ++k) n[k] = __ldg(np + k);
#pragma unroll
        for (int k = 0; k < R; ++k) v[k] = IO<T, V>::ldg(src + n[k] * C);
#pragma unroll
        for (int k = 0; k < R; ++k) IO<T, V>::stg(op + k * C, DIFF ? vsub<V>(v[k], c) : v[k]);
        op += R * C;
        np += R;
    }
}
}  // namespace

namespace pcs {
int launch_flat_rows(const float *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                     int64_t C, float *out, cudaStream_t s, const char *what);
int launch_flat_rows_bf16(const void *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                          int64_t C, void *out, cudaStream_t s, const char *what);
}

static int neighbor_scatter_impl(bool bf, const void *ifeats, const int32_t *nidxs, const int32_t *lens,
                                 const int32_t *bgs, const int32_t *cidxs, int64_t pn, int64_t ifn, int64_t csum,
                                 int use_diff, void *sfeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_scatter: negative size");
    if (pn == 0 || ifn == 0 || csum == 0) return 0;
    cudaStream_t s = as_stream(stream);
    const int es = bf ? 2 : 4;
    // Two kernels: the edge-parallel flat one (needs the centre of every edge, cidxs, for use_diff) and the
    // centre-walking one (centre row in registers, one index->gather chain per row in flight).  The walk wins for
    // wide rows; thresholds (in fp32-equivalent columns = row bytes / 4) are tunable (pcs_set_option).
    const bool vec16 = ((ifn * es) % 16 == 0) && aligned(ifeats, 16);
    const int64_t walk_min = use_diff ? (vec16 ? opt::scatter_walk_min_c_diff.load() : opt::scatter_walk_min_c_diff_odd.load())
                                      : (vec16 ? opt::scatter_walk_min_c_copy.load() : opt::scatter_walk_min_c_copy_odd.load());
    const bool walk = (ifn * es) / 4 >= walk_min;
    if (aligned(sfeats, 16) && !walk && (!use_diff || cidxs)) {
        if (bf) return launch_flat_rows_bf16(ifeats, pn, nidxs, use_diff ? cidxs : nullptr, csum, ifn, sfeats, s, "neighbor_scatter");
        return launch_flat_rows(static_cast<const float *>(ifeats), pn, nidxs, use_diff ? cidxs : nullptr, csum, ifn,
                                static_cast<float *>(sfeats), s, "neighbor_scatter");
    }
    RowTiling t = make_tiling_es(ifn, es, common_align({ifeats, sfeats}));
    int64_t grid = grid_for(pn, t, kBlock);
#define PCS_SCATTER_R(R_)                                                                                                          \
    do {                                                                                                                           \
        if (use_diff) k_neighbor_scatter<T, V, R_, true><<<grid, kBlock, 0, s>>>((const T *)ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, (T *)sfeats); \
        else k_neighbor_scatter<T, V, R_, false><<<grid, kBlock, 0, s>>>((const T *)ifeats, nidxs, lens, bgs, pn, ifn, t.gshift, t.CH, (T *)sfeats);        \
    } while (0)
    const int64_t rn = opt::walk_rows_narrow.load();
    PCS_DISPATCH_TV(bf, t.V,
        if (V * (int)sizeof(T) >= 16 || rn < 8) PCS_SCATTER_R(4); else PCS_SCATTER_R(8));
#undef PCS_SCATTER_R
    return check_launch("neighbor_scatter");
}

extern "C" int pcs_neighbor_scatter(const float *ifeats, const int32_t *nidxs, const int32_t *lens,
                                    const int32_t *bgs, const int32_t *cidxs, int64_t pn, int64_t ifn,
                                    int64_t csum, int use_diff, float *sfeats, pcs_stream_t stream) {
    return neighbor_scatter_impl(false, ifeats, nidxs, lens, bgs, cidxs, pn, ifn, csum, use_diff, sfeats, stream);
}
extern "C" int pcs_neighbor_scatter_bf16(const void *ifeats, const int32_t *nidxs, const int32_t *lens,
                                         const int32_t *bgs, const int32_t *cidxs, int64_t pn, int64_t ifn,
                                         int64_t csum, int use_diff, void *sfeats, pcs_stream_t stream) {
    return neighbor_scatter_impl(true, ifeats, nidxs, lens, bgs, cidxs, pn, ifn, csum, use_diff, sfeats, stream);
}

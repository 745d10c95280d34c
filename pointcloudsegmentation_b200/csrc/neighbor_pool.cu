// neighbor_pool.cu — row-routing operators: a4 SumFeatScatter, a6 MaxFeatScatter, a8 LocationWeightSumBackward,
// a15 PermutateFeature; fp32 or bf16 storage.  All pure streaming: lanes over channels, no
// memset + scattered-write pairs.  Narrow / odd rows go to the TMA-staged broadcast of seg_stream.cu.
#include "common.cuh"

using namespace pcs;

namespace pcs {
int launch_flat_rows(const float *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                     int64_t C, float *out, cudaStream_t s, const char *what);
int launch_flat_rows_bf16(const void *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                          int64_t C, void *out, cudaStream_t s, const char *what);
int launch_seg_stream_bcast(bool bf, bool maxs, const void *in, const int32_t *max_idxs, const int32_t *lens,
                            const int32_t *bgs, int64_t pn, int64_t C, int64_t n_out_rows, void *out, cudaStream_t s,
                            const char *what, bool *done);
}  // namespace pcs

namespace {
constexpr int kBlock = 256;
constexpr int kRows = 4;  // rows per group in the edge-parallel kernel

// out[r,:] = in[idx[r],:]   (a4: sumFeatScatter TFNeighborKernelNew.cu:32-47;
//                            a15: permutateFeatureKernel PoolingIndex/PermutateFeature.cu:5-19)
// only used when `out` is not 16-byte aligned (flat_rows.cu otherwise)
template <typename T, int V>
__global__ void __launch_bounds__(kBlock)
k_row_gather(const T *__restrict__ in, const int32_t *__restrict__ idx, int64_t n_rows, int64_t C,
             int gshift, int CH, T *__restrict__ out) {
    int64_t q; int col;
    const int64_t n_quads = (n_rows + kRows - 1) / kRows;
    if (!row_item(n_quads, gshift, CH, V, C, q, col)) return;
    const int64_t r0 = q * kRows;
    int64_t src[kRows];
#pragma unroll
    for (int k = 0; k < kRows; ++k) src[k] = (r0 + k < n_rows) ? (int64_t)idx[r0 + k] : -1;
    Vec<V> v[kRows];
#pragma unroll
    for (int k = 0; k < kRows; ++k)
        if (src[k] >= 0) v[k] = IO<T, V>::ldg(in + src[k] * C + col);
#pragma unroll
    for (int k = 0; k < kRows; ++k)
        if (src[k] >= 0) IO<T, V>::stg(out + (r0 + k) * C + col, v[k]);
}

// Zero-fill rows [r0, r1) of a [*, C] tensor for this lane's columns.
template <typename T, int V>
__device__ __forceinline__ void zero_rows(T *out, int64_t r0, int64_t r1, int64_t C, int col) {
    const Vec<V> z = vzero<V>();
    for (int64_t r = r0; r < r1; ++r) IO<T, V>::stg(out + r * C + col, z);
}

// a8 / a4: out[bg_i+j,:] = in[i,:]  (locWSumBackwardKernel TFNeighborKernel.cu:153-177).  Rows
// between consecutive segments (none when bgs is the scan of lens) and after the last segment get
// the zeros the reference's memset (:452) would have left there.
template <typename T, int V>
__global__ void __launch_bounds__(kBlock)
k_row_broadcast(const T *__restrict__ in, const int32_t *__restrict__ lens,
                const int32_t *__restrict__ bgs, int64_t pn, int64_t C, int64_t n_out_rows, int gshift,
                int CH, T *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const Vec<V> v = IO<T, V>::ldg(in + i * C + col);
    T *op = out + bg * C + col;
    for (int r = 0; r < len; ++r) { IO<T, V>::stg(op, v); op += C; }
    if (i == 0) zero_rows<T, V>(out, 0, bg, C, col);
    const int64_t next = (i + 1 < pn) ? (int64_t)bgs[i + 1] : n_out_rows;
    zero_rows<T, V>(out, bg + len, next, C, col);
}

// a6: sfeats[bg_i+r, f] = (max_idxs[i,f]==r) ? gfeats[i,f] : 0 — the memset + scattered 4-byte
// writes of maxFeatScatter (TFNeighborKernelNew.cu:121-139,204) become one coalesced pass.
template <typename T, int V>
__global__ void __launch_bounds__(kBlock)
k_max_scatter(const T *__restrict__ gfeats, const int32_t *__restrict__ max_idxs,
              const int32_t *__restrict__ bgs, const int32_t *__restrict__ lens, int64_t pn1,
              int64_t pn2, int64_t C, int gshift, int CH, T *__restrict__ out) {
    int64_t i; int col;
    if (!row_item(pn2, gshift, CH, V, C, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const Vec<V> g = IO<T, V>::ldg(gfeats + i * C + col);
    int mi[V];
#pragma unroll
    for (int k = 0; k < V; ++k) mi[k] = __ldg(max_idxs + i * C + col + k);
    T *op = out + bg * C + col;
    for (int r = 0; r < len; ++r) {
        Vec<V> o;
#pragma unroll
        for (int k = 0; k < V; ++k) o.v[k] = (mi[k] == r) ? g.v[k] : 0.f;
        IO<T, V>::stg(op, o);
        op += C;
    }
    if (i == 0) zero_rows<T, V>(out, 0, bg, C, col);
    const int64_t next = (i + 1 < pn2) ? (int64_t)bgs[i + 1] : pn1;
    zero_rows<T, V>(out, bg + len, next, C, col);
}
}  // namespace

static int launch_row_gather(bool bf, const void *in, int64_t in_rows, const int32_t *idx, int64_t n_rows, int64_t C,
                             void *out, cudaStream_t s, const char *what) {
    if (aligned(out, 16)) {
        if (bf) return launch_flat_rows_bf16(in, in_rows, idx, nullptr, n_rows, C, out, s, what);
        return launch_flat_rows(static_cast<const float *>(in), in_rows, idx, nullptr, n_rows, C, static_cast<float *>(out), s, what);
    }
    RowTiling t = make_tiling_es(C, bf ? 2 : 4, common_align({in, out}));
    int64_t grid = grid_for((n_rows + kRows - 1) / kRows, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, k_row_gather<T, V><<<grid, kBlock, 0, s>>>((const T *)in, idx, n_rows, C, t.gshift, t.CH, (T *)out));
    return check_launch(what);
}

static int row_broadcast(bool bf, const void *in, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                         int64_t n_out_rows, void *out, cudaStream_t s, const char *what) {
    bool done = false;
    if (launch_seg_stream_bcast(bf, false, in, nullptr, lens, bgs, pn, C, n_out_rows, out, s, what, &done)) return 1;
    if (done) return 0;
    RowTiling t = make_tiling_es(C, bf ? 2 : 4, common_align({in, out}));
    int64_t grid = grid_for(pn, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, k_row_broadcast<T, V><<<grid, kBlock, 0, s>>>((const T *)in, lens, bgs, pn, C, n_out_rows, t.gshift, t.CH, (T *)out));
    return check_launch(what);
}

static int sum_feat_scatter_impl(bool bf, const void *igfeats, const int32_t *cidxs, const int32_t *lens,
                                 const int32_t *bgs, int64_t pn, int64_t fd, int64_t csum, void *osfeats,
                                 pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fd >= 0 && csum >= 0, "neighbor_sum_feat_scatter: negative size");
    if (csum == 0 || fd == 0) return 0;  // reference early-out, TFNeighborKernelNew.cu:75
    cudaStream_t s = as_stream(stream);
    // With the CSR ranges at hand (the op's own nidxs_lens / nidxs_bgs inputs) this is a row broadcast:
    // read each centre row once, stream out its len copies; without them, a per-edge gather through cidxs.
    if (lens && bgs && pn > 0) return row_broadcast(bf, igfeats, lens, bgs, pn, fd, csum, osfeats, s, "neighbor_sum_feat_scatter");
    return launch_row_gather(bf, igfeats, pn, cidxs, csum, fd, osfeats, s, "neighbor_sum_feat_scatter");
}

extern "C" int pcs_neighbor_sum_feat_scatter(const float *igfeats, const int32_t *cidxs, const int32_t *lens,
                                             const int32_t *bgs, int64_t pn, int64_t fd, int64_t csum,
                                             float *osfeats, pcs_stream_t stream) {
    return sum_feat_scatter_impl(false, igfeats, cidxs, lens, bgs, pn, fd, csum, osfeats, stream);
}
extern "C" int pcs_neighbor_sum_feat_scatter_bf16(const void *igfeats, const int32_t *cidxs, const int32_t *lens,
                                                  const int32_t *bgs, int64_t pn, int64_t fd, int64_t csum,
                                                  void *osfeats, pcs_stream_t stream) {
    return sum_feat_scatter_impl(true, igfeats, cidxs, lens, bgs, pn, fd, csum, osfeats, stream);
}

// `elem_bytes` = 4 (fp32 / int32 rows, the reference's two instantiations) or 2 (bf16 rows)
static int permutate_impl(int elem_bytes, const void *feats, const int32_t *idxs, int64_t pn, int64_t ps, void *out,
                          pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ps >= 0, "permutate_feature: negative size");
    if (pn == 0 || ps == 0) return 0;
    return launch_row_gather(elem_bytes == 2, feats, pn, idxs, pn, ps, out, as_stream(stream), "permutate_feature");
}
extern "C" int pcs_permutate_feature(const void *feats, const int32_t *idxs, int64_t pn, int64_t ps,
                                     void *out, pcs_stream_t stream) {
    return permutate_impl(4, feats, idxs, pn, ps, out, stream);
}
extern "C" int pcs_permutate_feature_bf16(const void *feats, const int32_t *idxs, int64_t pn, int64_t ps,
                                          void *out, pcs_stream_t stream) {
    return permutate_impl(2, feats, idxs, pn, ps, out, stream);
}

static int locw_sum_backward_impl(bool bf, const void *dlw_sum, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                  int64_t m, int64_t csum, void *dlw, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && csum >= 0, "location_weight_sum_backward: negative size");
    if (csum == 0 || m == 0) return 0;
    cudaStream_t s = as_stream(stream);
    if (pn == 0) { PCS_CUDA(cudaMemsetAsync(dlw, 0, (bf ? 2 : 4) * csum * m, s)); return 0; }
    return row_broadcast(bf, dlw_sum, lens, bgs, pn, m, csum, dlw, s, "location_weight_sum_backward");
}
extern "C" int pcs_location_weight_sum_backward(const float *dlw_sum, const int32_t *lens, const int32_t *bgs,
                                                int64_t pn, int64_t m, int64_t csum, float *dlw,
                                                pcs_stream_t stream) {
    return locw_sum_backward_impl(false, dlw_sum, lens, bgs, pn, m, csum, dlw, stream);
}
extern "C" int pcs_location_weight_sum_backward_bf16(const void *dlw_sum, const int32_t *lens, const int32_t *bgs,
                                                     int64_t pn, int64_t m, int64_t csum, void *dlw,
                                                     pcs_stream_t stream) {
    return locw_sum_backward_impl(true, dlw_sum, lens, bgs, pn, m, csum, dlw, stream);
}

static int max_feat_scatter_impl(bool bf, const void *gfeats, const int32_t *max_idxs, const int32_t *vbegs,
                                 const int32_t *vlens, int64_t pn1, int64_t pn2, int64_t fd, void *sfeats,
                                 pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0 && fd >= 0, "neighbor_max_feat_scatter: negative size");
    if (pn1 == 0 || fd == 0) return 0;
    cudaStream_t s = as_stream(stream);
    const int es = bf ? 2 : 4;
    if (pn2 == 0) { PCS_CUDA(cudaMemsetAsync(sfeats, 0, es * pn1 * fd, s)); return 0; }
    bool done = false;
    if (launch_seg_stream_bcast(bf, true, gfeats, max_idxs, vlens, vbegs, pn2, fd, pn1, sfeats, s, "neighbor_max_feat_scatter", &done)) return 1;
    if (done) return 0;
    RowTiling t = make_tiling_es(fd, es, std::min(common_align({gfeats, sfeats}), 4 * es));
    int64_t grid = grid_for(pn2, t, kBlock);
    PCS_DISPATCH_TV(bf, t.V, k_max_scatter<T, V><<<grid, kBlock, 0, s>>>((const T *)gfeats, max_idxs, vbegs, vlens, pn1, pn2, fd, t.gshift, t.CH, (T *)sfeats));
    return check_launch("neighbor_max_feat_scatter");
}
extern "C" int pcs_neighbor_max_feat_scatter(const float *gfeats, const int32_t *max_idxs, const int32_t *vbegs,
                                             const int32_t *vlens, int64_t pn1, int64_t pn2, int64_t fd,
                                             float *sfeats, pcs_stream_t stream) {
    return max_feat_scatter_impl(false, gfeats, max_idxs, vbegs, vlens, pn1, pn2, fd, sfeats, stream);
}
extern "C" int pcs_neighbor_max_feat_scatter_bf16(const void *gfeats, const int32_t *max_idxs, const int32_t *vbegs,
                                                  const int32_t *vlens, int64_t pn1, int64_t pn2, int64_t fd,
                                                  void *sfeats, pcs_stream_t stream) {
    return max_feat_scatter_impl(true, gfeats, max_idxs, vbegs, vlens, pn1, pn2, fd, sfeats, stream);
}

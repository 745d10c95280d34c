// segment_stream.cu — streaming CSR segmented reductions: a3 NeighborSumFeatGather, a7
// LocationWeightSum, a9 LocationWeightFeatSum, a5 NeighborMaxFeatGather, a2 NeighborGather
// (SURVEY.md §8a).
//
// The INPUT is the big tensor ([csum, C] edge rows, read exactly once).  A group of G lanes (channel
// index on the lane axis, 16-byte lanes when C % 4 == 0) owns S consecutive centres and streams ALL
// their edge rows through a D-deep register ring: the row loads run D rows ahead of the accumulation
// and do not stop at segment boundaries, so short rows (degree 1..60) cost no extra latency — the
// per-centre metadata (lens/bgs) is read by a prefetch cursor ahead of use.  Rows are accumulated
// strictly in edge order per lane => deterministic sums, first-maximum argmax, no atomics, no memset
// (the reference: accumulator read-modify-written in global memory per step, fp32 atomicAdd for the
// gathers; sumFeatGather TFNeighborKernelNew.cu:4-27, maxFeatGather :83-118, featGather
// TFNeighborKernel.cu:63-125, locWSumForwardKernel :128-151, locationWeightFeatSumForwardKernel :179-211).
#include "common.cuh"
#include <stdlib.h>

using namespace pcs;

namespace {
constexpr int kBlock = 256;
constexpr int kD = 8;  // rows in flight per lane

enum { M_SUM = 0, M_WSUM = 1, M_MAX = 2, M_SUBSUM = 3 };

// Walks the (non-empty) segments of centres [c, c1): e = current edge, rem = edges left in segment c.
// The (len, start) of centre c+1 is fetched when centre c is entered, so crossing a segment boundary
// uses registers and the metadata latency hides behind the rows of the current segment.
struct Cursor {
    int64_t c, c1, e;
    int rem, nrem, ne;
    __device__ __forceinline__ void fetch_next(const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs) {
        nrem = 0; ne = 0;
        if (c + 1 < c1) { nrem = __ldg(lens + c + 1); ne = __ldg(bgs + c + 1); }
    }
    __device__ __forceinline__ void init(int64_t c0, int64_t cend, const int32_t *lens, const int32_t *bgs) {
        c = c0; c1 = cend; e = 0; rem = 0; nrem = 0; ne = 0;
        if (c < c1) {
            rem = __ldg(lens + c); e = __ldg(bgs + c);
            fetch_next(lens, bgs);
            if (rem == 0) advance(lens, bgs);
        }
    }
    __device__ __forceinline__ void advance(const int32_t *lens, const int32_t *bgs) {
        do {
            ++c;
            if (c >= c1) { rem = 0; return; }
            rem = nrem; e = ne;
            fetch_next(lens, bgs);
        } while (rem == 0);
    }
    __device__ __forceinline__ bool done() const { return c >= c1; }
    __device__ __forceinline__ void next(const int32_t *lens, const int32_t *bgs) {
        ++e;
        if (--rem == 0) advance(lens, bgs);
    }
};

// INDIRECT: the row of "edge position" p is feats[perm[p], :] (transpose CSR of NeighborGather);
// then lens/bgs describe positions in perm (bgs = tbgs, lens derived from tbgs by the caller kernel).
template <int V, int MODE, bool INDIRECT>
__global__ void __launch_bounds__(kBlock)
k_segment_stream(const float *__restrict__ feats, const float *__restrict__ lw, const int32_t *__restrict__ perm,
                 const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t n_out, int64_t C,
                 int64_t m, int64_t ofn, int gshift, int CH, int S, float *__restrict__ out,
                 int32_t *__restrict__ out_idx) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t gid = t >> gshift;
    const int lane = (int)(t & ((1 << gshift) - 1));
    const int64_t cb = gid / CH;
    const int ch = (int)(gid - cb * CH);
    const int col = ((ch << gshift) + lane) * V;
    const int64_t c0 = cb * S;
    if (c0 >= n_out || col >= C) return;
    const int64_t c1 = min(c0 + (int64_t)S, n_out);
    const int64_t wcol = (MODE == M_WSUM) ? col / ofn : 0;

    Cursor P, Q;  // P: prefetch (kD rows ahead), Q: consume
    P.init(c0, c1, lens, bgs);
    Q = P;
    Vec<V> buf[kD];
    float wbuf[kD];
#pragma unroll
    for (int k = 0; k < kD; ++k) {
        buf[k] = vzero<V>();
        wbuf[k] = 0.f;
        if (!P.done()) {
            const int64_t r = INDIRECT ? (int64_t)__ldg(perm + P.e) : P.e;
            buf[k] = ld<V>(feats + r * C + col);
            if (MODE == M_WSUM) wbuf[k] = __ldg(lw + r * m + wcol);
            P.next(lens, bgs);
        }
    }
    Vec<V> acc = vzero<V>();
    int bi[V];
#pragma unroll
    for (int j = 0; j < V; ++j) bi[j] = 0;
    int pos = 0;            // offset of the current row within its segment
    int64_t cw = c0;        // next centre whose output is still unwritten
    while (!Q.done()) {
#pragma unroll
        for (int k = 0; k < kD; ++k) {
            if (Q.done()) break;
            const Vec<V> v = buf[k];
            const float w = wbuf[k];
            if (!P.done()) {  // refill this ring slot before using v (independent of the math below)
                const int64_t r = INDIRECT ? (int64_t)__ldg(perm + P.e) : P.e;
                buf[k] = ld<V>(feats + r * C + col);
                if (MODE == M_WSUM) wbuf[k] = __ldg(lw + r * m + wcol);
                P.next(lens, bgs);
            }
            if (MODE == M_MAX) {
#pragma unroll
                for (int j = 0; j < V; ++j) {
                    if (pos == 0) { acc.v[j] = v.v[j]; bi[j] = 0; }
                    else if (acc.v[j] < v.v[j]) { acc.v[j] = v.v[j]; bi[j] = pos; }
                }
            } else if (MODE == M_WSUM) {
#pragma unroll
                for (int j = 0; j < V; ++j) acc.v[j] = fmaf(w, v.v[j], acc.v[j]);
            } else {
                vadd<V>(acc, v);
            }
            ++pos;
            const int64_t qc = Q.c;
            Q.next(lens, bgs);
            if (Q.c != qc) {  // segment qc finished: write it and the empty centres before it
                for (; cw < qc; ++cw) {
                    if (MODE != M_SUBSUM) st<V>(out + cw * C + col, vzero<V>());
                    if (MODE == M_MAX) st<V>(reinterpret_cast<float *>(out_idx) + cw * C + col, vzero<V>());
                }
                float *o = out + qc * C + col;
                if (MODE == M_SUBSUM) st<V>(o, vsub<V>(ld<V>(o), acc));
                else st<V>(o, acc);
                if (MODE == M_MAX) {
                    Vec<V> bv;
#pragma unroll
                    for (int j = 0; j < V; ++j) bv.v[j] = __int_as_float(bi[j]);
                    st<V>(reinterpret_cast<float *>(out_idx) + qc * C + col, bv);
                }
                cw = qc + 1;
                acc = vzero<V>();
                pos = 0;
            }
        }
    }
    for (; cw < c1; ++cw) {  // trailing empty centres
        if (MODE != M_SUBSUM) st<V>(out + cw * C + col, vzero<V>());
        if (MODE == M_MAX) st<V>(reinterpret_cast<float *>(out_idx) + cw * C + col, vzero<V>());
    }
}

__global__ void __launch_bounds__(kBlock)
k_lens_from_row_starts(const int32_t *__restrict__ tbgs, int64_t n, int32_t *__restrict__ lens) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) lens[i] = tbgs[i + 1] - tbgs[i];
}
}  // namespace

namespace pcs {
// Centres per group: stream ~128 edges per group, but keep >= ~4 waves of groups on 148 SMs.
static int pick_S(int64_t n_out, int64_t n_edges, int G, int CH) {
    const int64_t deg = n_out > 0 ? (n_edges + n_out - 1) / n_out : 1;
    int64_t S = 128 / (deg > 0 ? deg : 1);
    if (S < 1) S = 1;
    if (S > 64) S = 64;
    const int64_t want_groups = (int64_t)148 * 2048 * 4 / G;
    while (S > 1 && ((n_out + S - 1) / S) * CH < want_groups) S >>= 1;
    return (int)S;
}

template <int MODE, bool INDIRECT>
static int launch_stream(const float *feats, const float *lw, const int32_t *perm, const int32_t *lens,
                         const int32_t *bgs, int64_t n_out, int64_t n_edges, int64_t C, int64_t m, int64_t ofn,
                         float *out, int32_t *out_idx, cudaStream_t s, const char *what) {
    bool can4 = aligned(feats, 16) && aligned(out, 16) && (!out_idx || aligned(out_idx, 16)) &&
                (MODE != M_WSUM || ofn % 4 == 0);
    bool can2 = aligned(feats, 8) && aligned(out, 8) && (!out_idx || aligned(out_idx, 8)) &&
                (MODE != M_WSUM || ofn % 2 == 0);
    RowTiling t = make_tiling(C, can4, can2);
    const int G = 1 << t.gshift;
    const int S = pick_S(n_out, n_edges, G, t.CH);
    const int64_t groups = ((n_out + S - 1) / S) * t.CH;
    const int64_t grid = ((groups << t.gshift) + kBlock - 1) / kBlock;
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "%s: tensor too large", what);
    switch (t.V) {
        case 4: k_segment_stream<4, MODE, INDIRECT><<<(unsigned)grid, kBlock, 0, s>>>(feats, lw, perm, lens, bgs, n_out, C, m, ofn, t.gshift, t.CH, S, out, out_idx); break;
        case 2: k_segment_stream<2, MODE, INDIRECT><<<(unsigned)grid, kBlock, 0, s>>>(feats, lw, perm, lens, bgs, n_out, C, m, ofn, t.gshift, t.CH, S, out, out_idx); break;
        default: k_segment_stream<1, MODE, INDIRECT><<<(unsigned)grid, kBlock, 0, s>>>(feats, lw, perm, lens, bgs, n_out, C, m, ofn, t.gshift, t.CH, S, out, out_idx); break;
    }
    return check_launch(what);
}
}  // namespace pcs

namespace pcs {
int walk_segment_sum(const float *feats, const float *lw, const int32_t *lens, const int32_t *bgs, int64_t pn,
                     int64_t C, int64_t m, int64_t ofn, float *out, cudaStream_t s, const char *what);
int walk_neighbor_gather(const float *sfeats, const int32_t *lens, const int32_t *bgs, const int32_t *tperm,
                         const int32_t *tbgs, int64_t pn, int64_t ifn, int use_diff, float *ifeats, cudaStream_t s);
int walk_max_gather(const float *ifeats, const int32_t *vlens, const int32_t *vbegs, int64_t pn2, int64_t fd,
                    float *gfeats, int32_t *max_idxs, cudaStream_t s);
static bool use_walk() {
    static const int v = [] { const char *e = getenv("PCS_SEG_KERNEL"); return (e && e[0] == 'w') ? 1 : 0; }();
    return v != 0;
}
}  // namespace pcs

extern "C" int pcs_neighbor_sum_feat_gather(const float *ifeats, const int32_t *lens, const int32_t *bgs,
                                            int64_t pn, int64_t fd, int64_t csum, float *ogfeats,
                                            pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fd >= 0 && csum >= 0, "neighbor_sum_feat_gather: negative size");
    if (pn == 0 || fd == 0) return 0;
    if (use_walk()) return walk_segment_sum(ifeats, nullptr, lens, bgs, pn, fd, 0, 1, ogfeats, as_stream(stream), "neighbor_sum_feat_gather");
    return launch_stream<M_SUM, false>(ifeats, nullptr, nullptr, lens, bgs, pn, csum, fd, 0, 1, ogfeats, nullptr,
                                       as_stream(stream), "neighbor_sum_feat_gather");
}

extern "C" int pcs_location_weight_sum(const float *lw, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                       int64_t m, int64_t csum, float *lw_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && csum >= 0, "location_weight_sum: negative size");
    if (pn == 0 || m == 0) return 0;
    if (use_walk()) return walk_segment_sum(lw, nullptr, lens, bgs, pn, m, 0, 1, lw_sum, as_stream(stream), "location_weight_sum");
    return launch_stream<M_SUM, false>(lw, nullptr, nullptr, lens, bgs, pn, csum, m, 0, 1, lw_sum, nullptr,
                                       as_stream(stream), "location_weight_sum");
}

extern "C" int pcs_location_weight_feat_sum(const float *lw, const float *tfeats, const int32_t *lens,
                                            const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn,
                                            int64_t csum, float *otfeats_sum, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum: negative size");
    if (pn == 0 || m == 0 || ofn == 0) return 0;
    if (use_walk()) return walk_segment_sum(tfeats, lw, lens, bgs, pn, m * ofn, m, ofn, otfeats_sum, as_stream(stream), "location_weight_feat_sum");
    return launch_stream<M_WSUM, false>(tfeats, lw, nullptr, lens, bgs, pn, csum, m * ofn, m, ofn, otfeats_sum,
                                        nullptr, as_stream(stream), "location_weight_feat_sum");
}

extern "C" int pcs_neighbor_max_feat_gather(const float *ifeats, const int32_t *vlens, const int32_t *vbegs,
                                            int64_t pn1, int64_t pn2, int64_t fd, float *gfeats,
                                            int32_t *max_idxs, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0 && fd >= 0, "neighbor_max_feat_gather: negative size");
    if (pn2 == 0 || fd == 0) return 0;
    if (use_walk()) return walk_max_gather(ifeats, vlens, vbegs, pn2, fd, gfeats, max_idxs, as_stream(stream));
    return launch_stream<M_MAX, false>(ifeats, nullptr, nullptr, vlens, vbegs, pn2, pn1, fd, 0, 1, gfeats, max_idxs,
                                       as_stream(stream), "neighbor_max_feat_gather");
}

// a2: incoming edges through the transpose CSR (ascending edge id), then, for use_diff, minus the
// point's own segment sum (second streaming pass over the direct CSR).  tlens = workspace-free: the
// transpose row lengths live in the caller-provided tlens buffer (pn ints, from pcs_csr_transpose).
extern "C" int pcs_neighbor_gather(const float *sfeats, const int32_t *lens, const int32_t *bgs,
                                   const int32_t *tperm, const int32_t *tbgs, const int32_t *tlens, int64_t pn,
                                   int64_t ifn, int64_t csum, int use_diff, float *ifeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && ifn >= 0 && csum >= 0, "neighbor_gather: negative size");
    if (pn == 0 || ifn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    if (use_walk()) return walk_neighbor_gather(sfeats, lens, bgs, tperm, tbgs, pn, ifn, use_diff, ifeats, s);
    if (launch_stream<M_SUM, true>(sfeats, nullptr, tperm, tlens, tbgs, pn, csum, ifn, 0, 1, ifeats, nullptr, s,
                                   "neighbor_gather"))
        return 1;
    if (use_diff)
        return launch_stream<M_SUBSUM, false>(sfeats, nullptr, nullptr, lens, bgs, pn, csum, ifn, 0, 1, ifeats,
                                              nullptr, s, "neighbor_gather(diff)");
    return 0;
}

extern "C" int pcs_csr_row_lens(const int32_t *tbgs, int64_t pn, int32_t *tlens, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0, "csr_row_lens: negative size");
    if (pn == 0) return 0;
    k_lens_from_row_starts<<<(unsigned)((pn + kBlock - 1) / kBlock), kBlock, 0, as_stream(stream)>>>(tbgs, pn, tlens);
    return check_launch("csr_row_lens");
}

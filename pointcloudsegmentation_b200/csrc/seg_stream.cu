// seg_stream.cu — segmented sum / max over NARROW rows as a TMA-staged stream (a3 NeighborSumFeatGather,
// a7 LocationWeightSum, a5 NeighborMaxFeatGather at C*sizeof(T) below ~a cache line, and every odd width).
//
// Why: with rows of 12-140 bytes the walking kernels (segment_reduce.cu) cannot keep enough bytes in flight —
// a lane owns 4 bytes of a row, a warp's load touches a dozen partial sectors, and the kernel is bound by the
// L1 wavefront rate / load latency at 45-68 % of HBM (profiles/r3a_ops.log).  But the edge rows of CONSECUTIVE
// centres are one contiguous stream: [bgs[i0]*C, (bgs[i1-1]+lens[i1-1])*C).  So a CTA takes a tile of centres,
// has the TMA engine copy the tile's slice of the stream into shared memory in 16 KB bulk copies
// (cp.async.bulk + mbarrier complete_tx: no LSU instruction, no register staging, double-buffered), and every
// (centre, column-lane) item walks ITS rows out of shared memory with a running accumulator in registers —
// strictly in edge order, so the sum is the same sequential fp32 sum the reference kernel computes
// (sumFeatGather, TFNeighborKernelNew.cu:19-25) and the arg-max is the first maximum (:111).
// The output [pn, C] is written item-by-item = perfectly coalesced.
//
// Contiguity (bgs[i+1] == bgs[i] + lens[i]) is what every CSR of this code base satisfies, but the op's
// signature does not promise it: each CTA checks its own tile and falls back to walking global memory.
#include "common.cuh"
#include "tcgen05.cuh"

using namespace pcs;
using namespace pcs::tc;

namespace {
constexpr int kNT = 256;              // threads per CTA
// (centre, lane) items per thread: a tile is ~ items x mean degree x lane bytes of stream; narrow lanes get more items
// so that the tile prologue (CSR ranges -> first bulk copy: two dependent DRAM round trips) is amortised over
// >= 100 KB of stream whatever the lane width
template <int V> struct Ipt { static constexpr int value = V <= 2 ? 8 : 4; };
static inline int items_of(int V) { return kNT * (V <= 2 ? 8 : 4); }
constexpr int kMaxTc = 512;           // centres per tile (their bgs / lens live in shared memory)
constexpr int kChunkBytes = 16384;    // one bulk copy
constexpr int kStages = 2;

__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(bar) : "memory");
}

template <typename T, int V, bool MAX>
__global__ void __launch_bounds__(kNT)
k_seg_stream(const T *__restrict__ feats, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
             int64_t pn, int C, int Tc, int64_t total_elems, T *__restrict__ out, int32_t *__restrict__ max_idxs) {
    __shared__ __align__(128) unsigned char s_buf[kStages][kChunkBytes];
    __shared__ int s_bg[kMaxTc], s_len[kMaxTc];
    __shared__ __align__(8) uint64_t s_bar[kStages];
    constexpr int ES = (int)sizeof(T);
    constexpr int EPB = 16 / ES;                  // elements per 16 bytes
    constexpr uint32_t CHE = kChunkBytes / ES;    // elements per chunk
    constexpr int kIPT = Ipt<V>::value;
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * Tc;
    const int n_c = (int)min((int64_t)Tc, pn - i0);
    for (int t = tid; t < n_c; t += kNT) { s_bg[t] = bgs[i0 + t]; s_len[t] = lens[i0 + t]; }
    if (tid == 0) {
        mbar_init(smem_u32(&s_bar[0]), 1);
        mbar_init(smem_u32(&s_bar[1]), 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    bool ok = true;
    for (int t = tid; t < n_c; t += kNT) {
        ok = ok && s_len[t] >= 0 && s_bg[t] >= 0;
        if (t + 1 < n_c) ok = ok && (s_bg[t] + s_len[t] == s_bg[t + 1]);
    }
    const int CPI = C / V;  // lanes per centre
    const int n_items = n_c * CPI;
    const int64_t e_first = (int64_t)s_bg[0] * C;
    const int64_t e_last = ((int64_t)s_bg[n_c - 1] + s_len[n_c - 1]) * C;
    const int64_t a0 = e_first & ~(int64_t)(EPB - 1);  // tile start rounded down to 16 bytes (elements)
    ok = ok && (e_last <= total_elems) && (e_last - a0 < ((int64_t)1 << 31));
    ok = __syncthreads_and(ok);

    // per-item state: position of the next row (elements, relative to a0) and one past the last row
    uint32_t pos[kIPT], end[kIPT];
    float acc[kIPT][V];
    int bi[kIPT][V];
    int row[kIPT];
#pragma unroll
    for (int k = 0; k < kIPT; ++k) {
        const int item = tid + k * kNT;
        pos[k] = end[k] = 0;
        row[k] = 0;
#pragma unroll
        for (int q = 0; q < V; ++q) { acc[k][q] = 0.f; bi[k][q] = 0; }
        if (item < n_items) {
            const int ci = item / CPI;
            const int lane = item - ci * CPI;
            const int64_t p = (int64_t)s_bg[ci] * C + lane * V;
            if (ok) {
                pos[k] = (uint32_t)(p - a0);
                end[k] = pos[k] + (uint32_t)s_len[ci] * (uint32_t)C;
            } else {  // non-contiguous CSR: walk global memory (correctness path)
                const T *g = feats + p;
                for (int r = 0; r < s_len[ci]; ++r, g += C) {
                    const Vec<V> v = IO<T, V>::ldg(g);
#pragma unroll
                    for (int q = 0; q < V; ++q) {
                        if (MAX) { if (r == 0 || acc[k][q] < v.v[q]) { acc[k][q] = v.v[q]; bi[k][q] = r; } }
                        else acc[k][q] += v.v[q];
                    }
                }
            }
        }
    }

    if (ok) {
        const uint32_t span = (uint32_t)(e_last - a0);
        const uint32_t n_chunks = (span + CHE - 1) / CHE;
        // chunk q -> stage q & 1.  The bulk part is a multiple of 16 bytes that stays inside the tensor; the
        // (< 16 byte) remainder at the very end of the tensor is copied element-wise.
        auto issue = [&](uint32_t q) {
            const uint32_t c0 = q * CHE;
            const uint32_t n_el = min(CHE, span - c0);
            const int64_t abs0 = a0 + c0;
            const uint32_t want = (n_el * ES + 15u) & ~15u;
            const int64_t avail = (total_elems - abs0) * ES;
            const uint32_t bulk = (uint32_t)min((int64_t)want, avail & ~(int64_t)15);
            unsigned char *dst = s_buf[q & 1];
            if (tid == 0) {
                const uint32_t bar = smem_u32(&s_bar[q & 1]);
                if (bulk) { mbar_arrive_expect_tx(bar, bulk); bulk_g2s(dst, feats + abs0, bulk, bar); }
                else mbar_arrive(bar);
            }
            if (bulk < want) {
                const uint32_t b_el = bulk / ES;
                if (b_el + tid < n_el) reinterpret_cast<T *>(dst)[b_el + tid] = feats[abs0 + b_el + tid];
            }
        };
        if (n_chunks) issue(0);
        for (uint32_t q = 0; q < n_chunks; ++q) {
            if (q + 1 < n_chunks) issue(q + 1);  // its stage was released by the barrier that ended iteration q-1
            mbar_wait<0>(smem_u32(&s_bar[q & 1]), (q >> 1) & 1);  // stage q & 1 is on its (q >> 1)-th use
            __syncthreads();  // the element-wise tail (if any) is visible too
            const T *buf = reinterpret_cast<const T *>(s_buf[q & 1]);
            const uint32_t c0 = q * CHE;
            const uint32_t hi = c0 + min(CHE, span - c0);
#pragma unroll
            for (int k = 0; k < kIPT; ++k) {
                uint32_t p = pos[k];
                const uint32_t lim = min(end[k], hi);
                int r = row[k];
#pragma unroll 4
                for (; p < lim; p += C, ++r) {
                    const Vec<V> v = IO<T, V>::lds(buf + (p - c0));
#pragma unroll
                    for (int q2 = 0; q2 < V; ++q2) {
                        if (MAX) { if (r == 0 || acc[k][q2] < v.v[q2]) { acc[k][q2] = v.v[q2]; bi[k][q2] = r; } }
                        else acc[k][q2] += v.v[q2];
                    }
                }
                pos[k] = p;
                row[k] = r;
            }
            __syncthreads();  // everyone is done with this stage before the copy of chunk q+2 may land in it
        }
    }

    // out[(i0 + ci) * C + lane * V] = out[i0 * C + item * V]: consecutive items -> consecutive addresses
#pragma unroll
    for (int k = 0; k < kIPT; ++k) {
        const int item = tid + k * kNT;
        if (item < n_items) {
            Vec<V> o;
#pragma unroll
            for (int q = 0; q < V; ++q) o.v[q] = acc[k][q];
            const int64_t off = i0 * C + (int64_t)item * V;
            IO<T, V>::stg(out + off, o);
            if (MAX) {
                if constexpr (V == 1) max_idxs[off] = bi[k][0];
                else if constexpr (V == 2) *reinterpret_cast<int2 *>(max_idxs + off) = make_int2(bi[k][0], bi[k][1]);
                else {
#pragma unroll
                    for (int q = 0; q < V; q += 4)
                        *reinterpret_cast<int4 *>(max_idxs + off + q) = make_int4(bi[k][q], bi[k][q + 1], bi[k][q + 2], bi[k][q + 3]);
                }
            }
        }
    }
}

// ---- the mirror image: segment BROADCAST into a stream (a4 NeighborSumFeatScatter, a8 LocationWeightSumBackward,
// a6 NeighborMaxFeatScatter).  out[bg_i + r, :] = in[i, :]  (MAXS: g[i, f] where max_idxs[i, f] == r, else 0).
// Every (centre, lane) item keeps its V values in registers and writes them to ITS rows of the tile's output
// slice in shared memory; the slice leaves as 16 KB bulk stores (cp.async.bulk.global.shared::cta: no LSU store
// instruction, full lines whatever the row width), double-buffered.  The < 16-byte ragged ends of a tile's slice
// share their 16-byte piece with the neighbouring tile and go out element-wise.  A tile with a gap in its CSR
// (bgs[i+1] != bgs[i] + lens[i]) writes straight to global memory and zero-fills the gap rows, as the reference's
// memset would have (TFNeighborKernel.cu:452, TFNeighborKernelNew.cu:204).
template <typename T, int V, bool MAXS>
__global__ void __launch_bounds__(kNT)
k_seg_bcast(const T *__restrict__ in, const int32_t *__restrict__ max_idxs, const int32_t *__restrict__ lens,
            const int32_t *__restrict__ bgs, int64_t pn, int C, int Tc, int64_t n_out_rows, T *__restrict__ out) {
    __shared__ __align__(128) unsigned char s_buf[kStages][kChunkBytes];
    __shared__ int s_bg[kMaxTc + 1], s_len[kMaxTc];
    constexpr int ES = (int)sizeof(T);
    constexpr int EPB = 16 / ES;
    constexpr uint32_t CHE = kChunkBytes / ES;
    constexpr int kIPT = Ipt<V>::value;
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * Tc;
    const int n_c = (int)min((int64_t)Tc, pn - i0);
    for (int t = tid; t < n_c; t += kNT) { s_bg[t] = bgs[i0 + t]; s_len[t] = lens[i0 + t]; }
    if (tid == 0) s_bg[n_c] = (i0 + n_c < pn) ? bgs[i0 + n_c] : (int)min(n_out_rows, (int64_t)INT32_MAX);
    __syncthreads();
    bool ok = true;
    for (int t = tid; t < n_c; t += kNT) ok = ok && s_len[t] >= 0 && s_bg[t] >= 0 && (s_bg[t] + s_len[t] == s_bg[t + 1]);
    if (i0 == 0) ok = ok && s_bg[0] == 0;
    const int CPI = C / V;
    const int n_items = n_c * CPI;
    const int64_t e_first = (int64_t)s_bg[0] * C;
    const int64_t e_last = (int64_t)s_bg[n_c] * C;  // == end of the last segment when ok
    const int64_t a0 = e_first & ~(int64_t)(EPB - 1);
    ok = ok && (e_last <= n_out_rows * C) && (e_last - a0 < ((int64_t)1 << 31));
    ok = __syncthreads_and(ok);

    uint32_t pos[kIPT], end[kIPT];
    Vec<V> val[kIPT];
    int mi[kIPT][V];
    int row[kIPT];
#pragma unroll
    for (int k = 0; k < kIPT; ++k) {
        const int item = tid + k * kNT;
        pos[k] = end[k] = 0;
        row[k] = 0;
        val[k] = vzero<V>();
#pragma unroll
        for (int q = 0; q < V; ++q) mi[k][q] = -1;
        if (item < n_items) {
            const int ci = item / CPI;
            const int lane = item - ci * CPI;
            const int64_t src = (i0 + ci) * C + lane * V;
            val[k] = IO<T, V>::ldg(in + src);
            if (MAXS) {
#pragma unroll
                for (int q = 0; q < V; ++q) mi[k][q] = __ldg(max_idxs + src + q);
            }
            const int64_t p = (int64_t)s_bg[ci] * C + lane * V;
            if (ok) {
                pos[k] = (uint32_t)(p - a0);
                end[k] = pos[k] + (uint32_t)s_len[ci] * (uint32_t)C;
            } else {  // gap somewhere in this tile: direct stores + zero-filled gap rows
                T *g = out + p;
                const int len = s_len[ci];
                for (int r = 0; r < len; ++r, g += C) {
                    Vec<V> o = val[k];
                    if (MAXS) {
#pragma unroll
                        for (int q = 0; q < V; ++q) o.v[q] = (mi[k][q] == r) ? o.v[q] : 0.f;
                    }
                    IO<T, V>::stg(g, o);
                }
                const int64_t next = s_bg[ci + 1];
                const Vec<V> z = vzero<V>();
                for (int64_t r = (int64_t)s_bg[ci] + len; r < next; ++r) IO<T, V>::stg(out + r * C + lane * V, z);
                if (i0 + ci == 0)
                    for (int64_t r = 0; r < s_bg[0]; ++r) IO<T, V>::stg(out + r * C + lane * V, z);
            }
        }
    }

    if (ok && e_last > e_first) {
        const uint32_t span = (uint32_t)(e_last - a0);
        const uint32_t n_chunks = (span + CHE - 1) / CHE;
        const int64_t in_lo = (e_first + EPB - 1) & ~(int64_t)(EPB - 1);  // 16-byte pieces owned entirely by this tile
        const int64_t in_hi = e_last & ~(int64_t)(EPB - 1);
        for (uint32_t q = 0; q < n_chunks; ++q) {
            T *buf = reinterpret_cast<T *>(s_buf[q & 1]);
            // the bulk store issued two chunks ago read this stage: it must have finished READING shared memory
            if (tid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
            __syncthreads();
            const uint32_t c0 = q * CHE;
            const uint32_t hi = c0 + min(CHE, span - c0);
#pragma unroll
            for (int k = 0; k < kIPT; ++k) {
                uint32_t p = pos[k];
                const uint32_t lim = min(end[k], hi);
                int r = row[k];
#pragma unroll 4
                for (; p < lim; p += C, ++r) {
                    Vec<V> o = val[k];
                    if (MAXS) {
#pragma unroll
                        for (int q2 = 0; q2 < V; ++q2) o.v[q2] = (mi[k][q2] == r) ? o.v[q2] : 0.f;
                    }
                    IO<T, V>::stg(buf + (p - c0), o);
                }
                pos[k] = p;
                row[k] = r;
            }
            fence_async_smem();  // generic-proxy writes -> visible to the bulk (async-proxy) read
            __syncthreads();
            const int64_t lo = max(a0 + (int64_t)c0, in_lo), up = min(a0 + (int64_t)hi, in_hi);
            if (tid == 0 && up > lo) {
                asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                             ::"l"(out + lo), "r"(smem_u32(buf + (lo - a0 - c0))), "r"((uint32_t)((up - lo) * ES)) : "memory");
            }
            if (tid == 0) asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            // ragged ends (< 16 bytes each): element-wise
            if (q == 0) {
                const int64_t he = min(in_lo, e_last);
                if (e_first + tid < he) out[e_first + tid] = buf[e_first + tid - a0];
            }
            if (q + 1 == n_chunks) {
                const int64_t ts = max(in_hi, in_lo);
                if (ts + tid < e_last && ts >= a0 + (int64_t)c0) out[ts + tid] = buf[ts + tid - a0 - c0];
            }
        }
        if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // shared memory outlives the stores
    }
}

// pcs_set_option("seg_stream", v): 0 = never, 2 = whenever a tile fits, 1 (default) = where it measured faster than
// the walkers of segment_reduce.cu / neighbor_pool.cu.  The rule below is read off the full A/B on a B200
// (profiles/r3i_ab_stream_vs_walk.txt: every op x {f32, bf16} x C in {3, 12, 32, 35, 64, 128, 131}, m = 26):
//   sum reductions   stream for 4-byte-or-narrower lanes (odd widths) and rows up to 48 bytes; the walkers win once a
//                    lane is wider and the row is >= 52 bytes (C = 32/64/128 both dtypes, m = 26);
//   max reductions   stream only for 4-lane rows (f32 C % 4 == 0; bf16 C % 4 == 0 but not % 8): the arg-max epilogue
//                    of the 8-wide bf16 lanes and the 1-wide odd lanes is cheaper in the walker;
//   broadcasts       f32: rows below 128 bytes or lanes narrower than 16 bytes; bf16: rows up to 48 bytes;
//   arg-max routing  rows up to 48 bytes.
enum { KIND_SUM, KIND_MAX, KIND_BCAST, KIND_BCAST_MAX };
// (deg = mean segment length, 0 if unknown: on the towers' graphs, f32 rows of 128-256 bytes, the walker wins from a mean
// degree of ~18 up — 5.5 vs 4.2 TB/s at degree 33, 4.1 vs 4.4 at 14, tools/ab_max_gather.py)
static bool use_stream(int64_t C, int es, int V, int kind, int64_t deg = 0) {
    const int64_t mode = opt::seg_stream.load();
    if (mode <= 0 || C / V > items_of(V)) return false;
    if (mode >= 2) return true;
    const int64_t row = C * es;
    switch (kind) {
        case KIND_SUM: return V == 1 || row <= 48;
        case KIND_MAX: return V == 4 && !(es == 4 && row >= 128 && deg >= 18);
        case KIND_BCAST: return es == 4 ? (row < 128 || V * es < 16) : row <= 48;
        default: return row <= 48;
    }
}

template <bool MAX>
static int launch(bool bf, const void *feats, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                  int64_t rows_total, void *out, int32_t *max_idxs, cudaStream_t s, const char *what, bool *done) {
    *done = false;
    const int es = bf ? 2 : 4;
    if (!aligned(feats, 16) || C > (1 << 20)) return 0;
    int align = common_align({out});
    if (MAX) {  // V ints of arg-max go out in 16-byte pieces
        const int ia = common_align({max_idxs});
        const int vmax = ia >= 16 ? 8 : (ia >= 8 ? 2 : 1);
        align = std::min(align, vmax * es);
    }
    const RowTiling t = make_tiling_es(C, es, align);
    if (!use_stream(C, es, t.V, MAX ? KIND_MAX : KIND_SUM, pn > 0 ? rows_total / pn : 0)) return 0;
    const int CPI = (int)(C / t.V);
    const int Tc = std::max(1, std::min(kMaxTc, items_of(t.V) / CPI));
    const int64_t grid = (pn + Tc - 1) / Tc;
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "%s: too many tiles", what);
    PCS_DISPATCH_TV(bf, t.V,
        k_seg_stream<T, V, MAX><<<(unsigned)grid, kNT, 0, s>>>((const T *)feats, lens, bgs, pn, (int)C, Tc, rows_total * C,
                                                              (T *)out, max_idxs));
    *done = true;
    return check_launch(what);
}

static int launch_bcast(bool bf, bool maxs, const void *in, const int32_t *max_idxs, const int32_t *lens, const int32_t *bgs,
                        int64_t pn, int64_t C, int64_t n_out_rows, void *out, cudaStream_t s, const char *what, bool *done) {
    *done = false;
    const int es = bf ? 2 : 4;
    if (!aligned(out, 16) || C > (1 << 20) || n_out_rows >= INT32_MAX) return 0;
    int align = common_align({in});
    if (maxs) align = std::min(align, 4 * es);  // arg-max ints are read one by one; keep V <= 4
    const RowTiling t = make_tiling_es(C, es, align);
    if (!use_stream(C, es, t.V, maxs ? KIND_BCAST_MAX : KIND_BCAST)) return 0;
    const int CPI = (int)(C / t.V);
    const int Tc = std::max(1, std::min(kMaxTc, items_of(t.V) / CPI));
    const int64_t grid = (pn + Tc - 1) / Tc;
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "%s: too many tiles", what);
    PCS_DISPATCH_TV(bf, t.V,
        if (maxs) k_seg_bcast<T, V, true><<<(unsigned)grid, kNT, 0, s>>>((const T *)in, max_idxs, lens, bgs, pn, (int)C, Tc, n_out_rows, (T *)out);
        else k_seg_bcast<T, V, false><<<(unsigned)grid, kNT, 0, s>>>((const T *)in, max_idxs, lens, bgs, pn, (int)C, Tc, n_out_rows, (T *)out));
    *done = true;
    return check_launch(what);
}
}  // namespace

namespace pcs {
int launch_seg_stream_sum(bool bf, const void *feats, const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t C,
                          int64_t csum, void *out, cudaStream_t s, const char *what, bool *done) {
    return launch<false>(bf, feats, lens, bgs, pn, C, csum, out, nullptr, s, what, done);
}
int launch_seg_stream_max(bool bf, const void *feats, const int32_t *lens, const int32_t *bgs, int64_t pn2, int64_t C,
                          int64_t pn1, void *gfeats, int32_t *max_idxs, cudaStream_t s, const char *what, bool *done) {
    return launch<true>(bf, feats, lens, bgs, pn2, C, pn1, gfeats, max_idxs, s, what, done);
}
int launch_seg_stream_bcast(bool bf, bool maxs, const void *in, const int32_t *max_idxs, const int32_t *lens,
                            const int32_t *bgs, int64_t pn, int64_t C, int64_t n_out_rows, void *out, cudaStream_t s,
                            const char *what, bool *done) {
    return launch_bcast(bf, maxs, in, max_idxs, lens, bgs, pn, C, n_out_rows, out, s, what, done);
}
}  // namespace pcs

// flat_rows.cu — edge-parallel row routing: a1 NeighborScatter, a4 NeighborSumFeatScatter,
// a15 PermutateFeature.
//
// out[r, :] = in[src[r], :] (- in[cen[r], :] when DIFF).  The OUTPUT is the big tensor ([csum, C],
// k̄x the point tensor) and it is dense, so the kernel is indexed by the flat output: every thread
// owns whole 16-byte pieces of `out` regardless of C (C = 3, 35, 131 ... all store with STG.128, no
// idle lanes, no per-centre serial walk, perfect load balance over ragged degrees).  Source rows come
// from the small point tensor (mostly L2 hits): LDG.128 when C % 4 == 0, scalar otherwise.
// The reference walks each row serially in one thread with the point index on threadIdx.x
// (featScatter TFNeighborKernel.cu:34-59, sumFeatScatter TFNeighborKernelNew.cu:32-47,
// permutateFeatureKernel PoolingIndex/PermutateFeature.cu:5-19).
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;
constexpr int kU = 4;  // independent 16-byte pieces per thread (memory-level parallelism)

// C % 4 == 0: a piece never straddles a row.  IDX = uint32_t when everything fits 32 bits.
template <typename IDX, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_vec(const float *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                IDX n_pieces, IDX C4, float4 *__restrict__ out) {
    const IDX base = (IDX)blockIdx.x * (kBlock * kU) + threadIdx.x;
    IDX row[kU], c4[kU];
    int32_t s[kU], ce[kU];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        row[k] = i / C4;
        c4[k] = i - row[k] * C4;
        s[k] = 0; ce[k] = 0;
        if (i < n_pieces) {
            s[k] = __ldg(src + row[k]);
            if (DIFF) ce[k] = __ldg(cen + row[k]);
        }
    }
    float4 v[kU], w[kU];
    const float4 *in4 = reinterpret_cast<const float4 *>(in);
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            v[k] = __ldg(in4 + (int64_t)s[k] * C4 + c4[k]);
            if (DIFF) w[k] = __ldg(in4 + (int64_t)ce[k] * C4 + c4[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            float4 o = v[k];
            if (DIFF) { o.x = __fsub_rn(o.x, w[k].x); o.y = __fsub_rn(o.y, w[k].y); o.z = __fsub_rn(o.z, w[k].z); o.w = __fsub_rn(o.w, w[k].w); }
            __stcs(out + i, o);  // streaming store: the edge tensor is written once, keep L2 for the point rows
        }
    }
}

// any C: a 16-byte piece may straddle rows; 4 scalar gathers per piece, one vector store.
template <typename IDX, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_any(const float *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                IDX total, IDX C, float *__restrict__ out) {
    const IDX base = (IDX)blockIdx.x * (kBlock * kU) + threadIdx.x;
    float o[kU][4];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX f = (base + (IDX)k * kBlock) * 4;
        IDX row = f / C;
        IDX c = f - row * C;
        int32_t s = 0, ce = 0;
        if (f < total) { s = __ldg(src + row); if (DIFF) ce = __ldg(cen + row); }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            o[k][j] = 0.f;
            if (f + j < total) {
                float v = __ldg(in + (int64_t)s * C + c);
                if (DIFF) v = __fsub_rn(v, __ldg(in + (int64_t)ce * C + c));
                o[k][j] = v;
                if (++c == C) {
                    c = 0; ++row;
                    if (f + j + 1 < total) { s = __ldg(src + row); if (DIFF) ce = __ldg(cen + row); }
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX f = (base + (IDX)k * kBlock) * 4;
        if (f + 3 < total) {
            __stcs(reinterpret_cast<float4 *>(out + f), make_float4(o[k][0], o[k][1], o[k][2], o[k][3]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (f + j < total) out[f + j] = o[k][j];
        }
    }
}

// 32-bit fast path of the above for C >= 2 (a 16-byte piece then touches at most two rows): both row
// indices are fetched up front so the four gathers of a piece are independent (the generic kernel
// reloads the index in the middle of the piece: two serial index->gather chains), the division by C is
// a multiply-high, and addressing is 32-bit.  Requires total, pn*C < 2^31.
struct FastDiv {
    uint32_t d, mul, shr;
};
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv &f) {  // n < 2^31
    return (__umulhi(n, f.mul) + n) >> f.shr;
}

template <bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_two(const float *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                uint32_t total, uint32_t n_rows, FastDiv C, float *__restrict__ out) {
    const uint32_t base = blockIdx.x * (kBlock * kU) + threadIdx.x;
    uint32_t c0[kU], a0[kU], a1[kU], b0[kU], b1[kU];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * 4;
        a0[k] = a1[k] = b0[k] = b1[k] = 0; c0[k] = 0;
        if (f < total) {
            const uint32_t row = fdiv(f, C);
            c0[k] = f - row * C.d;
            const uint32_t row1 = min(row + 1, n_rows - 1);
            a0[k] = (uint32_t)__ldg(src + row) * C.d;
            a1[k] = (uint32_t)__ldg(src + row1) * C.d;
            if (DIFF) { b0[k] = (uint32_t)__ldg(cen + row) * C.d; b1[k] = (uint32_t)__ldg(cen + row1) * C.d; }
        }
    }
    float o[kU][4];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * 4;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            o[k][j] = 0.f;
            if (f + j < total) {
                const uint32_t c = c0[k] + j;
                const bool nxt = c >= C.d;
                const uint32_t cc = nxt ? c - C.d : c;
                float v = __ldg(in + (nxt ? a1[k] : a0[k]) + cc);
                if (DIFF) v = __fsub_rn(v, __ldg(in + (nxt ? b1[k] : b0[k]) + cc));
                o[k][j] = v;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * 4;
        if (f + 3 < total) {
            __stcs(reinterpret_cast<float4 *>(out + f), make_float4(o[k][0], o[k][1], o[k][2], o[k][3]));
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (f + j < total) out[f + j] = o[k][j];
        }
    }
}
}  // namespace

namespace pcs {
static FastDiv make_fastdiv(uint32_t d) {
    FastDiv f;
    f.d = d;
    uint32_t L = 0;
    while ((1ull << L) < d) ++L;
    f.mul = (uint32_t)((((uint64_t)1 << 32) * (((uint64_t)1 << L) - d)) / d + 1);
    f.shr = L;
    return f;
}

// out[n_rows, C] = in[src[r], :] (- in[cen[r], :] if cen), in [in_rows, C].  `out` must be 16-byte aligned.
int launch_flat_rows(const float *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                     int64_t C, float *out, cudaStream_t s, const char *what) {
    const int64_t total = n_rows * C;
    if (total == 0) return 0;
    const bool vec = (C % 4 == 0) && aligned(in, 16);
    const bool small = total < ((int64_t)1 << 31);
    const int64_t pieces = (total + 3) / 4;
    const int64_t grid = (pieces + kBlock * kU - 1) / (kBlock * kU);
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "%s: tensor too large", what);
#define PCS_FLAT(IDX, DIFF)                                                                                         \
    do {                                                                                                            \
        if (vec) k_flat_rows_vec<IDX, DIFF><<<(unsigned)grid, kBlock, 0, s>>>(in, src, cen, (IDX)pieces, (IDX)(C / 4), \
                                                                               reinterpret_cast<float4 *>(out));  \
        else k_flat_rows_any<IDX, DIFF><<<(unsigned)grid, kBlock, 0, s>>>(in, src, cen, (IDX)total, (IDX)C, out);     \
    } while (0)
    if (!vec && small && C >= 2 && in_rows * C < ((int64_t)1 << 31)) {
        const FastDiv fd = make_fastdiv((uint32_t)C);
        if (cen) k_flat_rows_two<true><<<(unsigned)grid, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
        else k_flat_rows_two<false><<<(unsigned)grid, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
    } else if (small) { if (cen) PCS_FLAT(uint32_t, true); else PCS_FLAT(uint32_t, false); }
    else       { if (cen) PCS_FLAT(int64_t, true); else PCS_FLAT(int64_t, false); }
#undef PCS_FLAT
    return check_launch(what);
}
}  // namespace pcs

// flat_rows.cu — edge-parallel row routing: a1 NeighborScatter, a4 NeighborSumFeatScatter,
// a15 PermutateFeature; fp32 or bf16 storage.
//
// out[r, :] = in[src[r], :] (- in[cen[r], :] when DIFF).  The OUTPUT is the big tensor ([csum, C],
// k̄x the point tensor) and it is dense, so the kernel is indexed by the flat output: every thread
// owns whole 16-byte pieces of `out` regardless of C (C = 3, 35, 131 ... all store with STG.128, no
// idle lanes, no per-centre serial walk, perfect load balance over ragged degrees).  Source rows come
// from the small point tensor (mostly L2 hits): LDG.128 when a row is a whole number of 16-byte pieces,
// element loads otherwise.  The reference walks each row serially in one thread with the point index on
// threadIdx.x (featScatter TFNeighborKernel.cu:34-59, sumFeatScatter TFNeighborKernelNew.cu:32-47,
// permutateFeatureKernel PoolingIndex/PermutateFeature.cu:5-19).
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;
constexpr int kU = 4;  // independent 16-byte pieces per thread (memory-level parallelism)

template <typename T> struct Piece;  // one 16-byte piece <-> EPB fp32 values
template <> struct Piece<float> {
    static constexpr int EPB = 4;
    static __device__ __forceinline__ void sub(uint4 &o, const uint4 &w) {
        o.x = __float_as_uint(__fsub_rn(__uint_as_float(o.x), __uint_as_float(w.x)));
        o.y = __float_as_uint(__fsub_rn(__uint_as_float(o.y), __uint_as_float(w.y)));
        o.z = __float_as_uint(__fsub_rn(__uint_as_float(o.z), __uint_as_float(w.z)));
        o.w = __float_as_uint(__fsub_rn(__uint_as_float(o.w), __uint_as_float(w.w)));
    }
    static __device__ __forceinline__ uint4 pack(const float (&v)[4]) {
        return make_uint4(__float_as_uint(v[0]), __float_as_uint(v[1]), __float_as_uint(v[2]), __float_as_uint(v[3]));
    }
};
template <> struct Piece<bf16> {
    static constexpr int EPB = 8;
    static __device__ __forceinline__ uint32_t sub2(uint32_t a, uint32_t b) {
        return bf_pack(__fsub_rn(bf_lo(a), bf_lo(b)), __fsub_rn(bf_hi(a), bf_hi(b)));
    }
    static __device__ __forceinline__ void sub(uint4 &o, const uint4 &w) {
        o.x = sub2(o.x, w.x); o.y = sub2(o.y, w.y); o.z = sub2(o.z, w.z); o.w = sub2(o.w, w.w);
    }
    static __device__ __forceinline__ uint4 pack(const float (&v)[8]) {
        return make_uint4(bf_pack(v[0], v[1]), bf_pack(v[2], v[3]), bf_pack(v[4], v[5]), bf_pack(v[6], v[7]));
    }
};

// A row is a whole number of pieces (C % EPB == 0): a piece never straddles a row.  IDX = uint32_t when
// everything fits 32 bits.  CP = pieces per row.
template <typename T, typename IDX, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_vec(const uint4 *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                IDX n_pieces, IDX CP, uint4 *__restrict__ out) {
    const IDX base = (IDX)blockIdx.x * (kBlock * kU) + threadIdx.x;
    IDX row[kU], cp[kU];
    int32_t s[kU], ce[kU];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        row[k] = i / CP;
        cp[k] = i - row[k] * CP;
        s[k] = 0; ce[k] = 0;
        if (i < n_pieces) {
            s[k] = __ldg(src + row[k]);
            if (DIFF) ce[k] = __ldg(cen + row[k]);
        }
    }
    uint4 v[kU], w[kU];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            v[k] = __ldg(in + (int64_t)s[k] * CP + cp[k]);
            if (DIFF) w[k] = __ldg(in + (int64_t)ce[k] * CP + cp[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            uint4 o = v[k];
            if (DIFF) Piece<T>::sub(o, w[k]);
            __stcs(out + i, o);  // streaming store: the edge tensor is written once, keep L2 for the point rows
        }
    }
}

// Any C (64-bit indices): a piece may straddle rows; EPB element gathers per piece, one vector store.
template <typename T, typename IDX, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_any(const T *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                IDX total, IDX C, T *__restrict__ out) {
    constexpr int EPB = Piece<T>::EPB;
    const IDX base = (IDX)blockIdx.x * (kBlock * kU) + threadIdx.x;
    for (int k = 0; k < kU; ++k) {
        const IDX f = (base + (IDX)k * kBlock) * EPB;
        if (f >= total) continue;
        IDX row = f / C;
        IDX c = f - row * C;
        int32_t s = __ldg(src + row), ce = DIFF ? __ldg(cen + row) : 0;
        float o[EPB];
#pragma unroll
        for (int j = 0; j < EPB; ++j) {
            o[j] = 0.f;
            if (f + j < total) {
                float v = IO<T, 1>::ldg(in + (int64_t)s * C + c).v[0];
                if (DIFF) v = __fsub_rn(v, IO<T, 1>::ldg(in + (int64_t)ce * C + c).v[0]);
                o[j] = v;
                if (++c == C) {
                    c = 0; ++row;
                    if (f + j + 1 < total) { s = __ldg(src + row); if (DIFF) ce = __ldg(cen + row); }
                }
            }
        }
        if (f + EPB - 1 < total) {
            __stcs(reinterpret_cast<uint4 *>(out + f), Piece<T>::pack(o));
        } else {
#pragma unroll
            for (int j = 0; j < EPB; ++j)
                if (f + j < total) { Vec<1> t; t.v[0] = o[j]; IO<T, 1>::stg(out + f + j, t); }
        }
    }
}

// 32-bit fast path of the above: a 16-byte piece touches at most NR rows (NR = 2 when C >= EPB - 1, up to 4 for
// bf16 rows of 3 elements); all NR row indices are fetched up front so the EPB gathers of a piece are independent
// (the generic kernel reloads the index in the middle of the piece: serial index->gather chains), the division by
// C is a multiply-high, and addressing is 32-bit.  Requires total, in_rows*C < 2^31.
struct FastDiv {
    uint32_t d, mul, shr;
};
__device__ __forceinline__ uint32_t fdiv(uint32_t n, const FastDiv &f) {  // n < 2^31
    return (__umulhi(n, f.mul) + n) >> f.shr;
}

template <typename T, int NR, bool DIFF>
__global__ void __launch_bounds__(kBlock)
k_flat_rows_two(const T *__restrict__ in, const int32_t *__restrict__ src, const int32_t *__restrict__ cen,
                uint32_t total, uint32_t n_rows, FastDiv C, T *__restrict__ out) {
    constexpr int EPB = Piece<T>::EPB;
    const uint32_t base = blockIdx.x * (kBlock * kU) + threadIdx.x;
    uint32_t c0[kU], a[kU][NR], b[kU][NR];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * EPB;
        c0[k] = 0;
#pragma unroll
        for (int r = 0; r < NR; ++r) a[k][r] = b[k][r] = 0;
        if (f < total) {
            const uint32_t row = fdiv(f, C);
            c0[k] = f - row * C.d;
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const uint32_t rr = min(row + r, n_rows - 1);
                a[k][r] = (uint32_t)__ldg(src + rr) * C.d;
                if (DIFF) b[k][r] = (uint32_t)__ldg(cen + rr) * C.d;
            }
        }
    }
    float o[kU][EPB];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * EPB;
#pragma unroll
        for (int j = 0; j < EPB; ++j) {
            o[k][j] = 0.f;
            if (f + j < total) {
                uint32_t c = c0[k] + j;
                uint32_t ra = a[k][0], rb = b[k][0];
#pragma unroll
                for (int r = 1; r < NR; ++r) {
                    const bool nxt = c >= C.d;
                    c = nxt ? c - C.d : c;
                    ra = nxt ? a[k][r] : ra;
                    if (DIFF) rb = nxt ? b[k][r] : rb;
                }  // r-th wrap -> slot r: wraps are consecutive, so the slot index is the wrap count
                float v = IO<T, 1>::ldg(in + ra + c).v[0];
                if (DIFF) v = __fsub_rn(v, IO<T, 1>::ldg(in + rb + c).v[0]);
                o[k][j] = v;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const uint32_t f = (base + k * kBlock) * EPB;
        if (f + EPB - 1 < total) {
            __stcs(reinterpret_cast<uint4 *>(out + f), Piece<T>::pack(o[k]));
        } else {
#pragma unroll
            for (int j = 0; j < EPB; ++j)
                if (f + j < total) { Vec<1> t; t.v[0] = o[k][j]; IO<T, 1>::stg(out + f + j, t); }
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

template <typename T>
static int launch_flat_rows_t(const T *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                              int64_t C, T *out, cudaStream_t s, const char *what) {
    constexpr int EPB = Piece<T>::EPB;
    const int64_t total = n_rows * C;
    if (total == 0) return 0;
    const bool vec = (C % EPB == 0) && aligned(in, 16);
    const bool small = total < ((int64_t)1 << 31);
    const int64_t pieces = (total + EPB - 1) / EPB;
    const int64_t grid = (pieces + kBlock * kU - 1) / (kBlock * kU);
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "%s: tensor too large", what);
    const unsigned g = (unsigned)grid;
    if (vec) {
        const uint4 *in4 = reinterpret_cast<const uint4 *>(in);
        uint4 *out4 = reinterpret_cast<uint4 *>(out);
        if (small) {
            if (cen) k_flat_rows_vec<T, uint32_t, true><<<g, kBlock, 0, s>>>(in4, src, cen, (uint32_t)pieces, (uint32_t)(C / EPB), out4);
            else k_flat_rows_vec<T, uint32_t, false><<<g, kBlock, 0, s>>>(in4, src, cen, (uint32_t)pieces, (uint32_t)(C / EPB), out4);
        } else {
            if (cen) k_flat_rows_vec<T, int64_t, true><<<g, kBlock, 0, s>>>(in4, src, cen, pieces, C / EPB, out4);
            else k_flat_rows_vec<T, int64_t, false><<<g, kBlock, 0, s>>>(in4, src, cen, pieces, C / EPB, out4);
        }
    } else if (small && in_rows * C < ((int64_t)1 << 31) && C >= 2) {
        const FastDiv fd = make_fastdiv((uint32_t)C);
        const bool two = C >= EPB - 1;  // a piece touches at most 2 rows
        if (two) {
            if (cen) k_flat_rows_two<T, 2, true><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
            else k_flat_rows_two<T, 2, false><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
        } else {
            if (cen) k_flat_rows_two<T, 4, true><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
            else k_flat_rows_two<T, 4, false><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)n_rows, fd, out);
        }
    } else if (small) {
        if (cen) k_flat_rows_any<T, uint32_t, true><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)C, out);
        else k_flat_rows_any<T, uint32_t, false><<<g, kBlock, 0, s>>>(in, src, cen, (uint32_t)total, (uint32_t)C, out);
    } else {
        if (cen) k_flat_rows_any<T, int64_t, true><<<g, kBlock, 0, s>>>(in, src, cen, total, C, out);
        else k_flat_rows_any<T, int64_t, false><<<g, kBlock, 0, s>>>(in, src, cen, total, C, out);
    }
    return check_launch(what);
}

// out[n_rows, C] = in[src[r], :] (- in[cen[r], :] if cen), in [in_rows, C].  `out` must be 16-byte aligned.
int launch_flat_rows(const float *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                     int64_t C, float *out, cudaStream_t s, const char *what) {
    return launch_flat_rows_t<float>(in, in_rows, src, cen, n_rows, C, out, s, what);
}
int launch_flat_rows_bf16(const void *in, int64_t in_rows, const int32_t *src, const int32_t *cen, int64_t n_rows,
                          int64_t C, void *out, cudaStream_t s, const char *what) {
    // a pure copy of rows with an even element count is a copy of fp32 words: half the element loads
    if (!cen && C % 2 == 0 && aligned(in, 4))
        return launch_flat_rows_t<float>(static_cast<const float *>(in), in_rows, src, nullptr, n_rows, C / 2,
                                         static_cast<float *>(out), s, what);
    return launch_flat_rows_t<bf16>(static_cast<const bf16 *>(in), in_rows, src, cen, n_rows, C,
                                    static_cast<bf16 *>(out), s, what);
}
}  // namespace pcs

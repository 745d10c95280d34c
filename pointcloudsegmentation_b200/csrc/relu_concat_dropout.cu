// relu_concat_dropout.cu — the glue between the classifier's linear layers (classifier_v3, model.py:783-793):
//     out = dropout(concat([relu(a), b], axis=1), keep_prob)
// in ONE pass over the rows, and its backward in one pass too.  As separate kernels (relu, cat, dropout and
// their three backward kernels) the 643- and 387-wide per-point tensors of a 64-block step cross HBM six times
// each way; here a and b are read once and out is written once.  No mask is stored: the backward regenerates
// the keep bits from (seed, row, column) with the same counter-based generator (Philox4x32-10), and takes the
// sign of relu from `out` (out > 0  <=>  kept and a > 0), which the next linear layer's weight gradient keeps
// alive anyway.
//
// Mapping (shared by forward and backward, it defines the mask): `out` is row-major contiguous [M, W = Na+Nb];
// a row is cut into G = ceil(W / 128) groups of 128 columns; a warp owns a row, lane l the columns
// 128 g + 32 j + l (j = 0..3) of group g, and draws their four 32-bit randoms from Philox counter
// (32 (row G + g) + l, 0), key = seed: random j decides column 128 g + 32 j + l.  An element is kept iff
// random < keep_prob * 2^32; kept values are scaled by 1 / keep_prob (tf.nn.dropout).  Every global access of a
// warp is a contiguous 128-byte row segment at a compile-time offset from the row pointer, and four groups
// (16 loads per lane, 2 KB per warp) are in flight at once.  (Two earlier versions walked the flat element
// index instead: one division and a wrap test per element made them instruction-bound at 3.1 TB/s.)
// tests/test_gpu_relu_concat_dropout.py compares with a numpy restatement of generator and mapping, bit for bit.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t k0, uint32_t k1) {
    uint32_t x0 = c0, x1 = c1, x2 = 0u, x3 = 0u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, x0), lo0 = 0xD2511F53u * x0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, x2), lo1 = 0xCD9E8D57u * x2;
        x0 = hi1 ^ x1 ^ k0;
        x1 = lo1;
        x2 = hi0 ^ x3 ^ k1;
        x3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return make_uint4(x0, x1, x2, x3);
}

struct Shape {
    int64_t M;
    int Na, Nb, W, G;
    uint32_t thresh, k0, k1;
    float scale;
};

constexpr int kU = 4;       // groups per pass
constexpr int kE = 4 * kU;  // columns per lane per pass

// keep bits of the lane's kE columns of groups g0 .. g0+kU-1 of `row` (bit 4u + j)
template <bool DROP>
__device__ __forceinline__ uint32_t keep_bits(const Shape &s, int64_t row, int g0, int lane) {
    if (!DROP) return 0xFFFFFFFFu;
    uint32_t bits = 0u;
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        if (g0 + u < s.G) {  // uniform
            const uint64_t ctr = ((uint64_t)row * (uint32_t)s.G + (uint32_t)(g0 + u)) * 32u + (uint32_t)lane;
            const uint4 r = philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), s.k0, s.k1);
            bits |= ((r.x < s.thresh ? 1u : 0u) | (r.y < s.thresh ? 2u : 0u) | (r.z < s.thresh ? 4u : 0u) | (r.w < s.thresh ? 8u : 0u)) << (4 * u);
        }
    }
    return bits;
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock)
k_rcd_forward(const float *__restrict__ a, int64_t lda, const float *__restrict__ b, int64_t ldb, Shape s,
              float *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (kBlock / 32);
    for (int64_t row = (int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5); row < s.M; row += warps) {
        const float *arow = a + row * lda;
        const float *brow = s.Nb ? b + row * ldb - s.Na : arow;  // indexed by the output column in both parts
        float *orow = out + row * s.W;
        for (int g0 = 0; g0 < s.G; g0 += kU) {
            const int c0 = g0 * 128 + lane;
            float v[kE];
            // all loads of the pass are in flight before the generator runs and before anything is stored
#pragma unroll
            for (int i = 0; i < kE; ++i) {
                const int col = c0 + 32 * i;
                v[i] = col < s.W ? __ldg((col < s.Na ? arow : brow) + col) : 0.f;
            }
            const uint32_t keep = keep_bits<DROP>(s, row, g0, lane);
#pragma unroll
            for (int i = 0; i < kE; ++i) {
                const int col = c0 + 32 * i;
                float x = col < s.Na ? fmaxf(v[i], 0.f) : v[i];
                if (DROP) x = ((keep >> i) & 1u) ? x * s.scale : 0.f;
                if (col < s.W) orow[col] = x;
            }
        }
    }
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock)
k_rcd_backward(const float *__restrict__ g, const float *__restrict__ out, Shape s, float *__restrict__ da,
               int64_t ldda, float *__restrict__ db, int64_t lddb) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (kBlock / 32);
    for (int64_t row = (int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5); row < s.M; row += warps) {
        const float *grow = g + row * s.W, *orow = out + row * s.W;
        float *darow = da + row * ldda;
        float *dbrow = db ? db + row * lddb - s.Na : nullptr;
        for (int g0 = 0; g0 < s.G; g0 += kU) {
            const int c0 = g0 * 128 + lane;
            float gv[kE], ov[kE];
#pragma unroll
            for (int i = 0; i < kE; ++i) {
                const int col = c0 + 32 * i;
                gv[i] = col < s.W ? __ldg(grow + col) : 0.f;
                ov[i] = col < s.Na ? __ldg(orow + col) : 0.f;
            }
            const uint32_t keep = keep_bits<DROP>(s, row, g0, lane);
#pragma unroll
            for (int i = 0; i < kE; ++i) {
                const int col = c0 + 32 * i;
                const float gs = DROP ? gv[i] * s.scale : gv[i];
                if (col < s.Na) darow[col] = ov[i] > 0.f ? gs : 0.f;  // out > 0 <=> kept and a > 0
                else if (col < s.W && dbrow) dbrow[col] = ((keep >> i) & 1u) ? gs : 0.f;
            }
        }
    }
}

int make_shape(int64_t M, int64_t Na, int64_t Nb, float keep_prob, uint64_t seed, Shape &s) {
    PCS_REQUIRE(M >= 0 && Na > 0 && Nb >= 0, "relu_concat_dropout: bad sizes M=%lld Na=%lld Nb=%lld", (long long)M, (long long)Na, (long long)Nb);
    PCS_REQUIRE(Na + Nb < (1 << 24) && M < (1ll << 40), "relu_concat_dropout: sizes out of range");
    PCS_REQUIRE(keep_prob > 0.f, "relu_concat_dropout: keep_prob must be in (0, 1] (>= 1 disables dropout)");
    s.M = M; s.Na = (int)Na; s.Nb = (int)Nb; s.W = (int)(Na + Nb);
    s.G = (s.W + 127) / 128;
    const double t = (double)keep_prob * 4294967296.0;
    s.thresh = t >= 4294967295.0 ? 4294967295u : (uint32_t)t;
    s.scale = 1.0f / keep_prob;
    s.k0 = (uint32_t)seed; s.k1 = (uint32_t)(seed >> 32);
    return 0;
}

int rcd_sm_count() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    }
    return n;
}

unsigned grid_for(const Shape &s) {
    const int64_t want = (s.M + (kBlock / 32) - 1) / (kBlock / 32);  // one row per warp and pass over the grid
    const int64_t cap = (int64_t)rcd_sm_count() * 8;                 // up to 8 resident CTAs of 256 threads per SM
    return (unsigned)(want < cap ? want : cap);
}
}  // namespace

extern "C" int pcs_relu_concat_dropout_forward(const float *a, int64_t M, int64_t Na, int64_t lda, const float *b, int64_t Nb,
                                               int64_t ldb, float keep_prob, uint64_t seed, float *out, pcs_stream_t stream) {
    Shape s;
    if (int rc = make_shape(M, Na, Nb, keep_prob, seed, s)) return rc;
    if (M == 0) return 0;
    PCS_REQUIRE(a && out && (b || Nb == 0), "relu_concat_dropout: null pointer");
    PCS_REQUIRE(lda >= Na && (Nb == 0 || ldb >= Nb), "relu_concat_dropout: row strides smaller than the row");
    cudaStream_t st = as_stream(stream);
    if (keep_prob < 1.f) k_rcd_forward<true><<<grid_for(s), kBlock, 0, st>>>(a, lda, b, ldb, s, out);
    else k_rcd_forward<false><<<grid_for(s), kBlock, 0, st>>>(a, lda, b, ldb, s, out);
    return check_launch("relu_concat_dropout_forward");
}

extern "C" int pcs_relu_concat_dropout_backward(const float *g, const float *out, int64_t M, int64_t Na, int64_t Nb, float keep_prob,
                                                uint64_t seed, float *da, int64_t ldda, float *db, int64_t lddb,
                                                pcs_stream_t stream) {
    Shape s;
    if (int rc = make_shape(M, Na, Nb, keep_prob, seed, s)) return rc;
    if (M == 0) return 0;
    PCS_REQUIRE(g && out && da, "relu_concat_dropout: null pointer");
    PCS_REQUIRE(ldda >= Na && (!db || lddb >= Nb), "relu_concat_dropout: row strides smaller than the row");
    cudaStream_t st = as_stream(stream);
    if (keep_prob < 1.f) k_rcd_backward<true><<<grid_for(s), kBlock, 0, st>>>(g, out, s, da, ldda, db, lddb);
    else k_rcd_backward<false><<<grid_for(s), kBlock, 0, st>>>(g, out, s, da, ldda, db, lddb);
    return check_launch("relu_concat_dropout_backward");
}

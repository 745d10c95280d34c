// relu_concat_dropout.cu — the glue between the classifier's linear layers (classifier_v3, model.py:783-793):
//     out = dropout(concat([relu(a), b], axis=1), keep_prob)
// in ONE pass over the rows, and its backward in one pass too.  As separate kernels (relu, cat, dropout and
// their three backward kernels) the 643- and 387-wide per-point tensors of a 64-block step cross HBM six times
// each way; here a and b are read once and out is written once.  No mask is stored: the backward regenerates
// the keep bits from (seed, element) with the same counter-based generator (Philox4x32-10), and takes the
// sign of relu from `out` (out > 0  <=>  kept and a > 0), which the next linear layer's weight gradient keeps
// alive anyway.
//
// Mapping (shared by forward and backward, it defines the mask): `out` is row-major contiguous [M, Na+Nb];
// its elements are cut into chunks of 128; lane l of the warp that owns chunk q handles elements
// 128 q + 32 j + l, j = 0..3, and draws its four 32-bit randoms from Philox counter (32 q + l, 0), key = seed:
// random j decides element j.  An element is kept iff random < keep_prob * 2^32; kept values are scaled by
// 1 / keep_prob (tf.nn.dropout).  Every global access of a warp is a contiguous 128-byte row segment.
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
    int64_t M, total, chunks;
    int Na, Nb, W;
    uint32_t thresh, k0, k1;
    float scale;
};

// the four elements of this lane in chunk q: element index, row, column (col < 0: past the end)
struct Quad {
    int64_t e[4], row[4];
    int col[4];
};

__device__ __forceinline__ Quad quad_of(const Shape &s, int64_t q, int lane) {
    Quad t;
    int64_t e = q * 128 + lane;
    int64_t row = s.total < (1ll << 32) ? (int64_t)((uint32_t)e / (uint32_t)s.W) : e / s.W;  // uniform branch
    int col = (int)(e - row * s.W);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        t.e[j] = e;
        t.row[j] = row;
        t.col[j] = e < s.total ? col : -1;
        e += 32;
        col += 32;
        while (col >= s.W) { col -= s.W; ++row; }
    }
    return t;
}

// keep bits of the four elements (bit j): Philox counter (32 q + lane, 0), key = seed
template <bool DROP>
__device__ __forceinline__ uint32_t keep_bits(const Shape &s, int64_t q, int lane) {
    if (!DROP) return 0xFu;
    const uint64_t ctr = (uint64_t)q * 32u + (uint64_t)lane;
    const uint4 r = philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), s.k0, s.k1);
    return (r.x < s.thresh ? 1u : 0u) | (r.y < s.thresh ? 2u : 0u) | (r.z < s.thresh ? 4u : 0u) | (r.w < s.thresh ? 8u : 0u);
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock)
k_rcd_forward(const float *__restrict__ a, int64_t lda, const float *__restrict__ b, int64_t ldb, Shape s,
              float *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (kBlock / 32);
    for (int64_t q = (int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5); q < s.chunks; q += warps) {
        const Quad t = quad_of(s, q, lane);
        float v[4];
        // all four loads are issued before the generator runs and before anything is stored
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float *src = t.col[j] < s.Na ? a + t.row[j] * lda + t.col[j] : b + t.row[j] * ldb + (t.col[j] - s.Na);
            v[j] = t.col[j] >= 0 ? __ldg(src) : 0.f;
        }
        const uint32_t keep = keep_bits<DROP>(s, q, lane);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float x = t.col[j] < s.Na ? fmaxf(v[j], 0.f) : v[j];
            if (DROP) x = ((keep >> j) & 1u) ? x * s.scale : 0.f;
            if (t.col[j] >= 0) out[t.e[j]] = x;
        }
    }
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock)
k_rcd_backward(const float *__restrict__ g, const float *__restrict__ out, Shape s, float *__restrict__ da,
               int64_t ldda, float *__restrict__ db, int64_t lddb) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = (int64_t)gridDim.x * (kBlock / 32);
    for (int64_t q = (int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5); q < s.chunks; q += warps) {
        const Quad t = quad_of(s, q, lane);
        float gv[4], ov[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            gv[j] = t.col[j] >= 0 ? __ldg(g + t.e[j]) : 0.f;
            ov[j] = (t.col[j] >= 0 && t.col[j] < s.Na) ? __ldg(out + t.e[j]) : 0.f;
        }
        const uint32_t keep = keep_bits<DROP>(s, q, lane);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (t.col[j] < 0) continue;
            const float gs = DROP ? gv[j] * s.scale : gv[j];
            if (t.col[j] < s.Na) da[t.row[j] * ldda + t.col[j]] = ov[j] > 0.f ? gs : 0.f;  // out > 0 <=> kept and a > 0
            else if (db) db[t.row[j] * lddb + (t.col[j] - s.Na)] = ((keep >> j) & 1u) ? gs : 0.f;
        }
    }
}

int make_shape(int64_t M, int64_t Na, int64_t Nb, float keep_prob, uint64_t seed, Shape &s) {
    PCS_REQUIRE(M >= 0 && Na > 0 && Nb >= 0, "relu_concat_dropout: bad sizes M=%lld Na=%lld Nb=%lld", (long long)M, (long long)Na, (long long)Nb);
    PCS_REQUIRE(Na + Nb < (1 << 24) && M < (1ll << 40), "relu_concat_dropout: sizes out of range");
    PCS_REQUIRE(keep_prob > 0.f, "relu_concat_dropout: keep_prob must be in (0, 1] (>= 1 disables dropout)");
    s.M = M; s.Na = (int)Na; s.Nb = (int)Nb; s.W = (int)(Na + Nb);
    s.total = M * (Na + Nb);
    s.chunks = (s.total + 127) / 128;
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
    const int64_t want = (s.chunks + (kBlock / 32) - 1) / (kBlock / 32);
    const int64_t cap = (int64_t)rcd_sm_count() * 8;  // 8 resident CTAs of 256 threads per SM
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

// relu_concat_dropout.cu — the glue between the classifier's linear layers (classifier_v3, model.py:783-793):
//     out = dropout(concat([relu(a), b], axis=1), keep_prob)
// in ONE pass over the rows, and its backward in one pass too.  As separate kernels (relu, cat, dropout and
// their three backward kernels) the 643- and 387-wide per-point tensors of a 64-block step cross HBM six times
// each way; here a and b are read once and out is written once.  No mask is stored: the backward regenerates
// the keep bits from (seed, element) with the same counter-based generator (Philox4x32-10), and takes the
// sign of relu from `out` (out > 0  <=>  kept and a > 0), which the next linear layer's weight gradient keeps
// alive anyway.
//
// Mask: `out` is row-major contiguous [M, W = Na+Nb]; element e = row W + col is word (e mod 128) / 32 of Philox
// counter (32 (e / 128) + e mod 32, 0) under key = seed; kept iff word < keep_prob * 2^32; kept values are scaled
// by 1 / keep_prob (tf.nn.dropout).
//
// Access pattern: both kernels walk the flat index space of `out` in chunks of 128 elements (lane l of a warp owns
// elements 128 q + 32 j + l): out (forward) and g, out (backward) move as full aligned 128-byte lines, the narrower
// operands (a, b / da, db) are touched at (row, col).  Tried and slower: one warp per row (every access misaligned),
// and a backward that walks da's and db's index spaces.  A warp iteration covers four chunks with all of its loads
// (2 KB per warp and operand) in flight; (row, col) come from one division per iteration and advance with 32-bit
// adds and one select (the first version divided and multiplied in 64 bits per element: instruction-bound at
// 3.1 TB/s).
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
    int Na, Nb, W;
    uint32_t thresh, k0, k1;
    float scale;
};

constexpr int kU = 4;       // chunks per warp iteration
constexpr int kE = 4 * kU;  // elements per lane and iteration

// Walk over the flat index space of a logical [M, N] matrix: position (row0 + .., col) plus two operand offsets
// relative to the operands' row-row0 pointers (row stride ld1 / ld2), all 32-bit (an iteration spans <= 512/N + 2 rows).
struct Steps {
    int N, adv_col, step1, step2, wrap1, wrap2;
};

struct Pos {
    int col, r1, r2;
};

__device__ __forceinline__ Steps steps_of(int N, int64_t ld1, int64_t ld2) {
    const int adv_row = 32 / N, adv_col = 32 % N;
    return {N, adv_col, adv_row * (int)ld1 + adv_col, adv_row * (int)ld2 + adv_col, (int)ld1 - N, (int)ld2 - N};
}

__device__ __forceinline__ void advance(const Steps &d, Pos &p) {
    p.col += d.adv_col;
    p.r1 += d.step1;
    p.r2 += d.step2;
    const bool wrap = p.col >= d.N;
    p.col -= wrap ? d.N : 0;
    p.r1 += wrap ? d.wrap1 : 0;
    p.r2 += wrap ? d.wrap2 : 0;
}

__device__ __forceinline__ int64_t row_of(int64_t e, int N, bool small) {
    return small ? (int64_t)((uint32_t)e / (uint32_t)N) : e / N;  // `small` is uniform
}

__device__ __forceinline__ uint32_t keep4(const uint4 &r, uint32_t thresh) {
    return (r.x < thresh ? 1u : 0u) | (r.y < thresh ? 2u : 0u) | (r.z < thresh ? 4u : 0u) | (r.w < thresh ? 8u : 0u);
}

template <bool DROP>
__device__ __forceinline__ uint32_t keep_bits(const Shape &s, int64_t q0, int lane) {
    if (!DROP) return 0xFFFFFFFFu;
    uint32_t keep = 0u;
#pragma unroll
    for (int u = 0; u < kU; ++u) {
        const uint64_t ctr = (uint64_t)(q0 + u) * 32u + (uint32_t)lane;
        keep |= keep4(philox4x32_10((uint32_t)ctr, (uint32_t)(ctr >> 32), s.k0, s.k1), s.thresh) << (4 * u);
    }
    return keep;
}

// One warp iteration of the forward.  TAIL = the 512 elements may reach past the end of the tensor (last warp only).
template <bool DROP, bool TAIL>
__device__ __forceinline__ void forward_iter(const float *__restrict__ a, int64_t lda, const float *__restrict__ b, int64_t ldb,
                                             const Shape &s, const Steps &d, float *__restrict__ out, int64_t q0, int lane,
                                             int64_t total, bool small) {
    const int64_t e0 = q0 * 128 + lane;
    const int left = TAIL ? (int)(total - e0 < 4096 ? total - e0 : 4096) : 4096;  // element i exists iff 32 i < left
    const int64_t row0 = row_of(e0 < total ? e0 : 0, s.W, small);
    Pos p;
    p.col = p.r1 = p.r2 = (int)(e0 - row0 * s.W);
    const float *a0 = a + row0 * lda;
    const float *b0 = s.Nb ? b + row0 * ldb - s.Na : a0;
    float v[kE];
    uint32_t relu_part = 0u;
    // all kE loads are in flight before the generator runs and before anything is stored; they are unconditional,
    // from a selected address: a conditional load (phi with a constant) lets the compiler fold the consumer into
    // the branch, which then waits for every load in turn
#pragma unroll
    for (int i = 0; i < kE; ++i) {
        const bool in_a = p.col < s.Na;
        relu_part |= (in_a ? 1u : 0u) << i;
        const float *src = in_a ? a0 + p.r1 : b0 + p.r2;
        v[i] = __ldg((!TAIL || 32 * i < left) ? src : a);
        advance(d, p);
    }
    const uint32_t keep = keep_bits<DROP>(s, q0, lane);
    float *dst = out + e0;
#pragma unroll
    for (int i = 0; i < kE; ++i) {
        float x = ((relu_part >> i) & 1u) ? fmaxf(v[i], 0.f) : v[i];
        if (DROP) x = ((keep >> i) & 1u) ? x * s.scale : 0.f;
        if (!TAIL || 32 * i < left) dst[32 * i] = x;
    }
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock)
k_rcd_forward(const float *__restrict__ a, int64_t lda, const float *__restrict__ b, int64_t ldb, Shape s,
              float *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t total = s.M * s.W, chunks = (total + 127) / 128, full = total / 128;
    const bool small = total < (1ll << 32);
    const int64_t step = (int64_t)gridDim.x * (kBlock / 32) * kU;
    const Steps d = steps_of(s.W, lda, ldb);
    for (int64_t q0 = ((int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5)) * kU; q0 < chunks; q0 += step) {
        if (q0 + kU <= full) forward_iter<DROP, false>(a, lda, b, ldb, s, d, out, q0, lane, total, small);
        else forward_iter<DROP, true>(a, lda, b, ldb, s, d, out, q0, lane, total, small);
    }
}

// Backward: the same walk over `out`'s index space (g and out are read as full aligned lines; da / db are written
// at (row, col)).  Walking da's and db's index spaces instead (aligned stores, gathered reads of the 643-wide
// operands) ran at half the speed.
template <bool DROP, bool TAIL>
__device__ __forceinline__ void backward_iter(const float *__restrict__ g, const float *__restrict__ out, const Shape &s,
                                              const Steps &d, float *__restrict__ da, int64_t ldda, float *__restrict__ db,
                                              int64_t lddb, int64_t q0, int lane, int64_t total, bool small) {
    const int64_t e0 = q0 * 128 + lane;
    const int left = TAIL ? (int)(total - e0 < 4096 ? total - e0 : 4096) : 4096;
    const int64_t row0 = row_of(e0 < total ? e0 : 0, s.W, small);
    Pos p;
    p.col = p.r1 = p.r2 = (int)(e0 - row0 * s.W);
    float *da0 = da + row0 * ldda;
    float *db0 = db ? db + row0 * lddb - s.Na : da0;  // da0: never stored through (see the store predicate)
    const float *gp = g + e0, *op = out + e0;
    float gv[kE], ov[kE];
    int r[kE];
    uint32_t relu_part = 0u;
#pragma unroll
    for (int i = 0; i < kE; ++i) {
        const bool in_a = p.col < s.Na;
        relu_part |= (in_a ? 1u : 0u) << i;
        const bool ok = !TAIL || 32 * i < left;  // loads are unconditional, from a selected address (see the forward)
        gv[i] = __ldg(ok ? gp + 32 * i : g);
        ov[i] = __ldg(ok && in_a ? op + 32 * i : out);
        r[i] = in_a ? p.r1 : p.r2;
        advance(d, p);
    }
    const uint32_t keep = keep_bits<DROP>(s, q0, lane);
#pragma unroll
    for (int i = 0; i < kE; ++i) {
        const float gs = DROP ? gv[i] * s.scale : gv[i];
        if (!TAIL || 32 * i < left) {
            if ((relu_part >> i) & 1u) da0[r[i]] = ov[i] > 0.f ? gs : 0.f;  // out > 0 <=> kept and a > 0
            else if (db) db0[r[i]] = ((keep >> i) & 1u) ? gs : 0.f;
        }
    }
}

template <bool DROP>
__global__ void __launch_bounds__(kBlock, 2)
k_rcd_backward(const float *__restrict__ g, const float *__restrict__ out, Shape s, float *__restrict__ da,
               int64_t ldda, float *__restrict__ db, int64_t lddb) {
    const int lane = threadIdx.x & 31;
    const int64_t total = s.M * s.W, chunks = (total + 127) / 128, full = total / 128;
    const bool small = total < (1ll << 32);
    const int64_t step = (int64_t)gridDim.x * (kBlock / 32) * kU;
    const Steps d = steps_of(s.W, ldda, lddb);
    for (int64_t q0 = ((int64_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5)) * kU; q0 < chunks; q0 += step) {
        if (q0 + kU <= full) backward_iter<DROP, false>(g, out, s, d, da, ldda, db, lddb, q0, lane, total, small);
        else backward_iter<DROP, true>(g, out, s, d, da, ldda, db, lddb, q0, lane, total, small);
    }
}

int make_shape(int64_t M, int64_t Na, int64_t Nb, float keep_prob, uint64_t seed, Shape &s) {
    PCS_REQUIRE(M >= 0 && Na > 0 && Nb >= 0, "relu_concat_dropout: bad sizes M=%lld Na=%lld Nb=%lld", (long long)M, (long long)Na, (long long)Nb);
    PCS_REQUIRE(Na + Nb < (1 << 20) && M < (1ll << 40), "relu_concat_dropout: sizes out of range");
    PCS_REQUIRE(keep_prob > 0.f, "relu_concat_dropout: keep_prob must be in (0, 1] (>= 1 disables dropout)");
    s.M = M; s.Na = (int)Na; s.Nb = (int)Nb; s.W = (int)(Na + Nb);
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

unsigned grid_for(int64_t elements) {
    const int64_t per_cta = (int64_t)(kBlock / 32) * kU * 128;
    const int64_t want = (elements + per_cta - 1) / per_cta + 1;
    const int64_t cap = (int64_t)rcd_sm_count() * 8;  // up to 8 resident CTAs of 256 threads per SM
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
    PCS_REQUIRE(lda < (1 << 20) && ldb < (1 << 20), "relu_concat_dropout: row strides out of range");
    cudaStream_t st = as_stream(stream);
    if (keep_prob < 1.f) k_rcd_forward<true><<<grid_for(M * (Na + Nb)), kBlock, 0, st>>>(a, lda, b, ldb, s, out);
    else k_rcd_forward<false><<<grid_for(M * (Na + Nb)), kBlock, 0, st>>>(a, lda, b, ldb, s, out);
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
    PCS_REQUIRE(ldda < (1 << 20) && lddb < (1 << 20), "relu_concat_dropout: row strides out of range");
    cudaStream_t st = as_stream(stream);
    if (keep_prob < 1.f) k_rcd_backward<true><<<grid_for(M * (Na + Nb)), kBlock, 0, st>>>(g, out, s, da, ldda, db, lddb);
    else k_rcd_backward<false><<<grid_for(M * (Na + Nb)), kBlock, 0, st>>>(g, out, s, da, ldda, db, lddb);
    return check_launch("relu_concat_dropout_backward");
}

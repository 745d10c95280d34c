// dense_mlp.cu — fused row-wise "dense-concat" MLP, the edge MLP of the reference's pointnet_conv /
// pointnet_pool / mlp (model_pointnet.py:10-80) after the split evaluation of network.py
// ("fused edge-conv layers").  No reference CUDA counterpart: upstream runs it as a
// chain of tf.contrib fully_connected + concat + relu over [csum, 2C+3+...] edge tensors.
//
// Per row r (one thread per row; rows are edges, ~10^6 of them):
//     u = [u1 (A) | u2 (B) | u3 (C) | uo (O)]                 pre-activations from the point-level GEMMs
//     h1 = relu(u1)
//     h2 = relu(u2 + W2 h1)                 W2 [B, A]
//     h3 = relu(u3 + W3 [h2, h1])           W3 [C, B+A]
//     y  = uo + Wo [h3, h2, h1]             Wo [O, C+B+A]      (concat order of the reference: newest first)
// Everything between u and y lives in registers; the three small weight matrices sit in shared memory
// and are read as warp-wide broadcasts (LDS.128 feeds 4 FFMA).  fp32 FFMA throughout: this work is
// FLOP-bound on the CUDA cores (1.3-5.4 kMAC per 0.4-0.8 KB row); tensor cores would need TF32/BF16
// operands, which the fp32 parity bar (1e-5) rules out.
// Backward recomputes h1..h3 from u (cheaper than storing them), and emits
//     du = [dpre1 | dpre2 | dpre3 | dy]  and  hcat = [h3 | h2 | h1]
// so that the weight gradients are three tall-skinny GEMMs on the host side.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kThreads = 128;

template <int N>
__device__ __forceinline__ void load_row(const float *__restrict__ p, float *dst) {
    static_assert(N % 4 == 0, "row pieces are multiples of 4 floats");
#pragma unroll
    for (int k = 0; k < N; k += 4) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(p + k));
        dst[k] = v.x; dst[k + 1] = v.y; dst[k + 2] = v.z; dst[k + 3] = v.w;
    }
}
template <int N>
__device__ __forceinline__ void store_row(float *__restrict__ p, const float *src) {
#pragma unroll
    for (int k = 0; k < N; k += 4)
        *reinterpret_cast<float4 *>(p + k) = make_float4(src[k], src[k + 1], src[k + 2], src[k + 3]);
}

// acc[k] += sum_j W[k][j] * x[j], W row-major [K][J] in shared memory (broadcast reads)
template <int K, int J>
__device__ __forceinline__ void matvec(const float *__restrict__ W, const float *x, float *acc) {
#pragma unroll
    for (int k = 0; k < K; ++k) {
        float a = acc[k];
#pragma unroll
        for (int j = 0; j < J; j += 4) {
            const float4 w = *reinterpret_cast<const float4 *>(W + k * J + j);
            a = fmaf(w.x, x[j], a); a = fmaf(w.y, x[j + 1], a); a = fmaf(w.z, x[j + 2], a); a = fmaf(w.w, x[j + 3], a);
        }
        acc[k] = a;
    }
}
template <int A, int B, int C, int O>
struct Smem {
    static_assert(A + B + C <= 64, "relu mask is one 64-bit word");
    float w2[B * A];
    float w3[C * (B + A)];
    float wo[O * (C + B + A)];
};

template <int A, int B, int C, int O>
__device__ __forceinline__ void load_weights(Smem<A, B, C, O> &s, const float *w2, const float *w3, const float *wo) {
    for (int i = threadIdx.x; i < B * A; i += kThreads) s.w2[i] = __ldg(w2 + i);
    for (int i = threadIdx.x; i < C * (B + A); i += kThreads) s.w3[i] = __ldg(w3 + i);
    for (int i = threadIdx.x; i < O * (C + B + A); i += kThreads) s.wo[i] = __ldg(wo + i);
    __syncthreads();
}

// h = [h3 (C) | h2 (B) | h1 (A)] from the row's pre-activations
template <int A, int B, int C, int O>
__device__ __forceinline__ void hidden(const Smem<A, B, C, O> &s, const float *__restrict__ ur, float *h) {
    float t[(C > B ? C : B) > A ? (C > B ? C : B) : A];
    load_row<A>(ur, t);
#pragma unroll
    for (int j = 0; j < A; ++j) h[C + B + j] = fmaxf(t[j], 0.f);
    load_row<B>(ur + A, t);
    matvec<B, A>(s.w2, h + C + B, t);
#pragma unroll
    for (int k = 0; k < B; ++k) h[C + k] = fmaxf(t[k], 0.f);
    load_row<C>(ur + A + B, t);
    matvec<C, B + A>(s.w3, h + C, t);
#pragma unroll
    for (int k = 0; k < C; ++k) h[k] = fmaxf(t[k], 0.f);
}

template <int A, int B, int C, int O>
__global__ void __launch_bounds__(kThreads)
k_dense_mlp3_fwd(const float *__restrict__ u, const float *__restrict__ w2, const float *__restrict__ w3,
                 const float *__restrict__ wo, int64_t rows, float *__restrict__ y) {
    __shared__ __align__(16) Smem<A, B, C, O> s;
    load_weights(s, w2, w3, wo);
    constexpr int H = A + B + C + O, HC = A + B + C;
    for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < rows; r += (int64_t)gridDim.x * kThreads) {
        const float *ur = u + r * H;
        float h[HC];
        hidden<A, B, C, O>(s, ur, h);
#pragma unroll
        for (int f0 = 0; f0 < O; f0 += 16) {  // 16 outputs at a time keeps the register count flat
            float acc[16];
            load_row<16>(ur + HC + f0, acc);
            matvec<16, HC>(s.wo + f0 * HC, h, acc);
            store_row<16>(y + r * O + f0, acc);
        }
    }
}

// Backward keeps TRANSPOSED copies of the weights in shared memory so that every dX product is again a
// "one accumulator, broadcast weight row" matvec (the scatter form  dx[j] += W[k][j] * g[k]  lets the
// scheduler hoist hundreds of weight loads and spills; measured with -Xptxas -v).
template <int A, int B, int C, int O>
struct SmemBwd {
    float w2[B * A];
    float w3[C * (B + A)];
    float w2t[A * B];            // [A][B]
    float w3t[(B + A) * C];      // [B+A][C]
    float wot[(C + B + A) * O];  // [HC][O]
};

//
// POOLED: the MLP output went straight into the segmented max (a5 NeighborMaxFeatGather), so dy is
// never materialised: row r of centre i = cidxs[r] gets dy[r,f] = dpool[i,f] if midx[i,f] == r - bgs[i]
// else 0 (exactly what a6 NeighborMaxFeatScatter would have written), read from the small, cache-resident
// [pn, O] tensors.
template <int A, int B, int C, int O, bool POOLED>
__global__ void __launch_bounds__(kThreads)
k_dense_mlp3_bwd(const float *__restrict__ u, const float *__restrict__ dy, const int32_t *__restrict__ midx,
                 const int32_t *__restrict__ cidxs, const int32_t *__restrict__ bgs, const float *__restrict__ w2,
                 const float *__restrict__ w3, const float *__restrict__ wo, int64_t rows,
                 float *__restrict__ du, float *__restrict__ hcat) {
    constexpr int H = A + B + C + O, HC = A + B + C;
    __shared__ __align__(16) SmemBwd<A, B, C, O> s;
    for (int i = threadIdx.x; i < B * A; i += kThreads) { const float v = __ldg(w2 + i); s.w2[i] = v; s.w2t[(i % A) * B + i / A] = v; }
    for (int i = threadIdx.x; i < C * (B + A); i += kThreads) { const float v = __ldg(w3 + i); s.w3[i] = v; s.w3t[(i % (B + A)) * C + i / (B + A)] = v; }
    for (int i = threadIdx.x; i < O * HC; i += kThreads) s.wot[(i % HC) * O + i / HC] = __ldg(wo + i);
    __syncthreads();
    for (int64_t r = (int64_t)blockIdx.x * kThreads + threadIdx.x; r < rows; r += (int64_t)gridDim.x * kThreads) {
        const float *ur = u + r * H;
        unsigned long long pos = 0ull;  // bit j = (h[j] > 0): all the backward needs of h once hcat is stored
        {
            float h[HC], t[(C > B ? C : B) > A ? (C > B ? C : B) : A];
            load_row<A>(ur, t);
#pragma unroll
            for (int j = 0; j < A; ++j) h[C + B + j] = fmaxf(t[j], 0.f);
            load_row<B>(ur + A, t);
            matvec<B, A>(s.w2, h + C + B, t);
#pragma unroll
            for (int k = 0; k < B; ++k) h[C + k] = fmaxf(t[k], 0.f);
            load_row<C>(ur + A + B, t);
            matvec<C, B + A>(s.w3, h + C, t);
#pragma unroll
            for (int k = 0; k < C; ++k) h[k] = fmaxf(t[k], 0.f);
            store_row<HC>(hcat + r * HC, h);
#pragma unroll
            for (int j = 0; j < HC; ++j) pos |= (h[j] > 0.f) ? (1ull << j) : 0ull;
        }
        float g[O];
        if (POOLED) {
            const int32_t i = __ldg(cidxs + r);
            const int32_t k = (int32_t)(r - __ldg(bgs + i));
            const float4 *gp = reinterpret_cast<const float4 *>(dy + (int64_t)i * O);
            const int4 *mp = reinterpret_cast<const int4 *>(midx + (int64_t)i * O);
#pragma unroll
            for (int f = 0; f < O; f += 4) {
                const float4 v = __ldg(gp + f / 4);
                const int4 m = __ldg(mp + f / 4);
                g[f] = m.x == k ? v.x : 0.f; g[f + 1] = m.y == k ? v.y : 0.f;
                g[f + 2] = m.z == k ? v.z : 0.f; g[f + 3] = m.w == k ? v.w : 0.f;
            }
        } else {
            load_row<O>(dy + r * O, g);
        }
        store_row<O>(du + r * H + HC, g);  // d(uo) = dy
        float d3[C], d2[B], d1[A];
#pragma unroll
        for (int k = 0; k < C; ++k) d3[k] = 0.f;
        matvec<C, O>(s.wot, g, d3);                              // dh3 = Wo[:, h3]^T dy
#pragma unroll
        for (int k = 0; k < C; ++k) d3[k] = ((pos >> k) & 1ull) ? d3[k] : 0.f;
        store_row<C>(du + r * H + A + B, d3);
#pragma unroll
        for (int k = 0; k < B; ++k) d2[k] = 0.f;
        matvec<B, O>(s.wot + C * O, g, d2);                      // dh2 = Wo[:, h2]^T dy + W3[:, h2]^T dpre3
        matvec<B, C>(s.w3t, d3, d2);
#pragma unroll
        for (int k = 0; k < B; ++k) d2[k] = ((pos >> (C + k)) & 1ull) ? d2[k] : 0.f;
        store_row<B>(du + r * H + A, d2);
#pragma unroll
        for (int j = 0; j < A; ++j) d1[j] = 0.f;
        matvec<A, O>(s.wot + (C + B) * O, g, d1);                // dh1 = Wo^T dy + W3^T dpre3 + W2^T dpre2
        matvec<A, C>(s.w3t + B * C, d3, d1);
        matvec<A, B>(s.w2t, d2, d1);
#pragma unroll
        for (int j = 0; j < A; ++j) d1[j] = ((pos >> (C + B + j)) & 1ull) ? d1[j] : 0.f;
        store_row<A>(du + r * H, d1);
    }
}

struct PoolRef {  // non-null midx selects the POOLED backward
    const int32_t *midx, *cidxs, *bgs;
};

template <int A, int B, int C, int O>
int launch(const float *u, const float *dy, PoolRef p, const float *w2, const float *w3, const float *wo, int64_t rows,
           float *y, float *du, float *hcat, cudaStream_t s) {
    int64_t grid = (rows + kThreads - 1) / kThreads;
    const int64_t cap = 148 * 16;  // grid-stride: the weights are staged once per CTA
    if (grid > cap) grid = cap;
    if (du && p.midx)
        k_dense_mlp3_bwd<A, B, C, O, true><<<(unsigned)grid, kThreads, 0, s>>>(u, dy, p.midx, p.cidxs, p.bgs, w2, w3, wo,
                                                                               rows, du, hcat);
    else if (du)
        k_dense_mlp3_bwd<A, B, C, O, false><<<(unsigned)grid, kThreads, 0, s>>>(u, dy, nullptr, nullptr, nullptr, w2, w3,
                                                                                wo, rows, du, hcat);
    else k_dense_mlp3_fwd<A, B, C, O><<<(unsigned)grid, kThreads, 0, s>>>(u, w2, w3, wo, rows, y);
    return check_launch(du ? "dense_mlp3_backward" : "dense_mlp3_forward");
}

int dispatch(int64_t A, int64_t B, int64_t C, int64_t O, const float *u, const float *dy, PoolRef p, const float *w2,
             const float *w3, const float *wo, int64_t rows, float *y, float *du, float *hcat, cudaStream_t s) {
#define PCS_CFG(a, b, c, o) \
    if (A == a && B == b && C == c && O == o) return launch<a, b, c, o>(u, dy, p, w2, w3, wo, rows, y, du, hcat, s);
    PCS_CFG(8, 8, 16, 32)
    PCS_CFG(16, 16, 32, 64)
    PCS_CFG(16, 16, 16, 48)
#undef PCS_CFG
    return fail("dense_mlp3: unsupported layer widths (%lld,%lld,%lld,%lld)", (long long)A, (long long)B, (long long)C,
                (long long)O);
}
}  // namespace

namespace pcs {
int dense_mlp3_forward_mma(const float *u, const float *w2, const float *w3, const float *wo, int64_t rows, int64_t A,
                           int64_t B, int64_t C, int64_t O, float *y, cudaStream_t s);
int dense_mlp3_backward_mma(const float *u, const float *dy, const int32_t *midx, const int32_t *cidxs,
                            const int32_t *bgs, const float *w2, const float *w3, const float *wo, int64_t rows,
                            int64_t A, int64_t B, int64_t C, int64_t O, float *du, float *hcat, float *wpart,
                            cudaStream_t s);
int64_t dense_mlp3_wgrad_partials_mma(int64_t rows, int64_t H);
}

extern "C" int pcs_dense_mlp3_supported(int64_t A, int64_t B, int64_t C, int64_t O) {
    return (A == 8 && B == 8 && C == 16 && O == 32) || (A == 16 && B == 16 && C == 32 && O == 64) ||
           (A == 16 && B == 16 && C == 16 && O == 48);
}

extern "C" int pcs_dense_mlp3_forward(const float *u, const float *w2, const float *w3, const float *wo,
                                      int64_t rows, int64_t A, int64_t B, int64_t C, int64_t O, float *y,
                                      pcs_stream_t stream) {
    PCS_REQUIRE(rows >= 0, "dense_mlp3_forward: negative size");
    PCS_REQUIRE(aligned(u, 16) && aligned(y, 16), "dense_mlp3_forward: tensors must be 16-byte aligned");
    if (rows == 0) return 0;
    if (opt::mlp_tensor_cores.load() && aligned(u, 8) && aligned(y, 8))
        return dense_mlp3_forward_mma(u, w2, w3, wo, rows, A, B, C, O, y, as_stream(stream));
    return dispatch(A, B, C, O, u, nullptr, PoolRef{nullptr, nullptr, nullptr}, w2, w3, wo, rows, y, nullptr, nullptr,
                    as_stream(stream));
}

extern "C" int pcs_dense_mlp3_backward(const float *u, const float *dy, const float *w2, const float *w3,
                                       const float *wo, int64_t rows, int64_t A, int64_t B, int64_t C, int64_t O,
                                       float *du, float *hcat, pcs_stream_t stream) {
    PCS_REQUIRE(rows >= 0, "dense_mlp3_backward: negative size");
    if (rows == 0) return 0;
    PCS_REQUIRE(du && hcat, "dense_mlp3_backward: outputs required");
    PCS_REQUIRE(aligned(u, 16) && aligned(dy, 16) && aligned(du, 16) && aligned(hcat, 16),
                "dense_mlp3_backward: tensors must be 16-byte aligned");
    if (rows == 0) return 0;
    if (opt::mlp_tensor_cores.load() && pcs_dense_mlp3_supported(A, B, C, O))
        return dense_mlp3_backward_mma(u, dy, nullptr, nullptr, nullptr, w2, w3, wo, rows, A, B, C, O, du, hcat, nullptr,
                                       as_stream(stream));
    return dispatch(A, B, C, O, u, dy, PoolRef{nullptr, nullptr, nullptr}, w2, w3, wo, rows, nullptr, du, hcat,
                    as_stream(stream));
}

extern "C" int64_t pcs_dense_mlp3_wgrad_partials(int64_t rows, int64_t A, int64_t B, int64_t C, int64_t O) {
    if (!opt::mlp_tensor_cores.load() || !pcs_dense_mlp3_supported(A, B, C, O)) return 0;
    return dense_mlp3_wgrad_partials_mma(rows, A + B + C + O);
}

extern "C" int pcs_dense_mlp3_backward_pooled(const float *u, const float *dpool, const int32_t *max_idxs,
                                              const int32_t *cidxs, const int32_t *nidxs_bgs, const float *w2,
                                              const float *w3, const float *wo, int64_t rows, int64_t A, int64_t B,
                                              int64_t C, int64_t O, float *du, float *hcat, float *wpart,
                                              pcs_stream_t stream) {
    PCS_REQUIRE(rows >= 0, "dense_mlp3_backward_pooled: negative size");
    if (rows == 0) return 0;
    PCS_REQUIRE(du && hcat, "dense_mlp3_backward_pooled: outputs required");
    PCS_REQUIRE(aligned(u, 16) && aligned(dpool, 16) && aligned(max_idxs, 16) && aligned(du, 16) && aligned(hcat, 16),
                "dense_mlp3_backward_pooled: tensors must be 16-byte aligned");
    if (rows == 0) return 0;
    PCS_REQUIRE(dpool && max_idxs && cidxs && nidxs_bgs, "dense_mlp3_backward_pooled: null pointer");
    if (opt::mlp_tensor_cores.load() && pcs_dense_mlp3_supported(A, B, C, O))
        return dense_mlp3_backward_mma(u, dpool, max_idxs, cidxs, nidxs_bgs, w2, w3, wo, rows, A, B, C, O, du, hcat,
                                       wpart, as_stream(stream));
    PCS_REQUIRE(wpart == nullptr, "dense_mlp3_backward_pooled: in-kernel weight gradients need mlp_tensor_cores=1");
    return dispatch(A, B, C, O, u, dpool, PoolRef{max_idxs, cidxs, nidxs_bgs}, w2, w3, wo, rows, nullptr, du, hcat,
                    as_stream(stream));
}

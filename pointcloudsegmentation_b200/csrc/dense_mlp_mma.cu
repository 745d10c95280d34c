// dense_mlp_mma.cu — the dense-concat edge MLP of dense_mlp.cu on the tensor cores.
//
// Why: the thread-per-row FFMA kernels read every weight as a shared-memory broadcast, and the shared
// memory return path (128 B/clk/SM) delivers one 4-byte operand per warp-FFMA: ncu shows l1tex at
// 84-94 % and the FMA pipe at 20-29 % (profiles/r1o_ncu_full_fused_raw.csv) — they are LDS-bound at a
// quarter of the fp32 peak.  A warp-level MMA reads one 8-byte weight fragment per 1024 MACs.
//
// How fp32 accuracy is kept: error-compensated TF32 ("3xTF32").  x = hi + lo with hi = tf32(x),
// lo = tf32(x - hi) (22 mantissa bits together), and  a*b ~= a_lo*b_hi + a_hi*b_lo + a_hi*b_hi  with fp32
// accumulation in the tensor core: ~1e-6 relative, inside the 1e-5 parity bar (tests/test_gpu_fused_ops.py
// checks against float64).
//
// Tiling: one warp owns 16 rows (edges) x all columns; `mma.sync.m16n8k8.tf32` (the warp-level path:
// K is only 8-64 here and the A operand is produced in registers by the previous layer, which tcgen05's
// shared-memory/TMEM operands would force through a round trip).  The D fragment of one layer IS the A
// fragment of the next without any shuffle: lane (g,t) holds D[g][2t], D[g][2t+1], D[g+8][2t],
// D[g+8][2t+1]; read as A[g][t], A[g][t+4], A[g+8][t], A[g+8][t+4] this is the same tile with the K index
// permuted (k<4 -> column 2k, k>=4 -> column 2(k-4)+1), and the matching B fragment (b0,b1) =
// (W[n][8kb+2t], W[n][8kb+2t+1]) is a plain float2 of the row-major weight matrix.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;

// fp32 -> tf32, round to nearest with ties away from zero (what cvt.rna.tf32.f32 computes for finite
// inputs) as two integer instructions: add half a tf32 ulp to the magnitude, clear the 13 low mantissa
// bits.  ptxas expands the cvt into ~8 instructions (NaN/Inf handling); with five splits per 8-column
// block that was most of the kernel's instruction stream (profiles/README.md, r1u source page: HMMA
// 10 % of warp instructions).  Non-finite inputs stay non-finite (Inf may become NaN), as they would in
// any product.
__device__ __forceinline__ uint32_t to_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }

// 8 consecutive columns of a 16-row tile as an MMA A operand, split into hi + lo
struct Frag {
    uint32_t hi[4], lo[4];
};
// c = {row g: col 2t, col 2t+1; row g+8: col 2t, col 2t+1}  ->  A order {a0,a1,a2,a3} = {c0,c2,c1,c3}
__device__ __forceinline__ Frag make_frag(float c0, float c1, float c2, float c3) {
    Frag f;
    const float v[4] = {c0, c2, c1, c3};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f.hi[i] = to_tf32(v[i]);
        f.lo[i] = to_tf32(v[i] - __uint_as_float(f.hi[i]));
    }
    return f;
}

__device__ __forceinline__ void mma_tf32(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// acc += a * W-fragment: small terms first.  wh / wl point at this lane's float2 of the hi / lo weight copy.
__device__ __forceinline__ void mma3(float (&acc)[4], const Frag &a, const float *wh, const float *wl) {
    const uint2 bh = *reinterpret_cast<const uint2 *>(wh);
    const uint2 bl = *reinterpret_cast<const uint2 *>(wl);
    mma_tf32(acc, a.lo, bh.x, bh.y);
    mma_tf32(acc, a.hi, bl.x, bl.y);
    mma_tf32(acc, a.hi, bh.x, bh.y);
}

// shared-memory row stride for a K-wide weight matrix: == 8 (mod 32) words, so the 4 rows x 8 words a
// half-warp reads with one LDS.64 fall into distinct banks
__host__ __device__ constexpr int padk(int K) { return ((K - 8 + 31) / 32) * 32 + 8; }

template <int A, int B, int C, int O>
struct Layout {
    static constexpr int K2 = padk(A), K3 = padk(A + B), KO = padk(A + B + C);
    static constexpr int o_w2 = 0, o_w3 = o_w2 + B * K2, o_wo = o_w3 + C * K3, n_half = o_wo + O * KO;
    static constexpr size_t bytes = 2 * sizeof(float) * n_half;  // hi copy then lo copy
};

// stage W [N][K] row-major as tf32 hi / lo with padded stride
__device__ __forceinline__ void stage(const float *__restrict__ w, int N, int K, int KP, float *hi, float *lo) {
    for (int i = threadIdx.x; i < N * K; i += kThreads) {
        const int n = i / K, k = i - n * K;
        const float v = __ldg(w + i);
        const uint32_t h = to_tf32(v);
        hi[n * KP + k] = __uint_as_float(h);
        lo[n * KP + k] = __uint_as_float(to_tf32(v - __uint_as_float(h)));
    }
}

template <int A, int B, int C, int O>
__global__ void __launch_bounds__(kThreads)  // (forcing more resident CTAs than the 106 / 156 / 192 registers allow: 9 % slower)
k_dense_mlp3_fwd_mma(const float *__restrict__ u, const float *__restrict__ w2, const float *__restrict__ w3,
                     const float *__restrict__ wo, int64_t rows, float *__restrict__ y) {
    using L = Layout<A, B, C, O>;
    constexpr int H = A + B + C + O, HC = A + B + C;
    extern __shared__ __align__(16) float smem[];
    float *hi = smem, *lo = smem + L::n_half;
    stage(w2, B, A, L::K2, hi + L::o_w2, lo + L::o_w2);
    stage(w3, C, A + B, L::K3, hi + L::o_w3, lo + L::o_w3);
    stage(wo, O, HC, L::KO, hi + L::o_wo, lo + L::o_wo);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int64_t n_tiles = (rows + 15) / 16;
    // software pipeline: the pre-activations of the warp's next tile are in flight while this one computes
    const int64_t stride = (int64_t)gridDim.x * kWarps;
    const int64_t first = (int64_t)blockIdx.x * kWarps + warp;
    float2 nu0[H / 8], nu1[H / 8];
    auto load_tile = [&](int64_t tile) {
        const int64_t r0 = tile * 16 + g, r1 = r0 + 8;
        const float *u0 = u + (r0 < rows ? r0 : rows - 1) * H + 2 * t;
        const float *u1 = u + (r1 < rows ? r1 : rows - 1) * H + 2 * t;
#pragma unroll
        for (int b = 0; b < H / 8; ++b) {
            nu0[b] = __ldg(reinterpret_cast<const float2 *>(u0 + 8 * b));
            nu1[b] = __ldg(reinterpret_cast<const float2 *>(u1 + 8 * b));
        }
    };
    if (first < n_tiles) load_tile(first);
    for (int64_t tile = first; tile < n_tiles; tile += stride) {
        const int64_t r0 = tile * 16 + g, r1 = r0 + 8;
        // cu[b] = 8-column block b of the tile in D layout
        float2 cu0[H / 8], cu1[H / 8];
#pragma unroll
        for (int b = 0; b < H / 8; ++b) { cu0[b] = nu0[b]; cu1[b] = nu1[b]; }
        if (tile + stride < n_tiles) load_tile(tile + stride);
        Frag h1[A / 8], h2[B / 8], h3[C / 8];
#pragma unroll
        for (int b = 0; b < A / 8; ++b)
            h1[b] = make_frag(fmaxf(cu0[b].x, 0.f), fmaxf(cu0[b].y, 0.f), fmaxf(cu1[b].x, 0.f), fmaxf(cu1[b].y, 0.f));
#pragma unroll
        for (int nt = 0; nt < B / 8; ++nt) {  // h2 = relu(u2 + W2 h1)
            float acc[4] = {cu0[A / 8 + nt].x, cu0[A / 8 + nt].y, cu1[A / 8 + nt].x, cu1[A / 8 + nt].y};
            const int w = L::o_w2 + (nt * 8 + g) * L::K2 + 2 * t;
#pragma unroll
            for (int kb = 0; kb < A / 8; ++kb) mma3(acc, h1[kb], hi + w + 8 * kb, lo + w + 8 * kb);
            h2[nt] = make_frag(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        }
#pragma unroll
        for (int nt = 0; nt < C / 8; ++nt) {  // h3 = relu(u3 + W3 [h2, h1])
            float acc[4] = {cu0[(A + B) / 8 + nt].x, cu0[(A + B) / 8 + nt].y, cu1[(A + B) / 8 + nt].x,
                            cu1[(A + B) / 8 + nt].y};
            const int w = L::o_w3 + (nt * 8 + g) * L::K3 + 2 * t;
#pragma unroll
            for (int kb = 0; kb < B / 8; ++kb) mma3(acc, h2[kb], hi + w + 8 * kb, lo + w + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < A / 8; ++kb) mma3(acc, h1[kb], hi + w + B + 8 * kb, lo + w + B + 8 * kb);
            h3[nt] = make_frag(fmaxf(acc[0], 0.f), fmaxf(acc[1], 0.f), fmaxf(acc[2], 0.f), fmaxf(acc[3], 0.f));
        }
#pragma unroll
        for (int nt = 0; nt < O / 8; ++nt) {  // y = uo + Wo [h3, h2, h1]
            float acc[4] = {cu0[HC / 8 + nt].x, cu0[HC / 8 + nt].y, cu1[HC / 8 + nt].x, cu1[HC / 8 + nt].y};
            const int w = L::o_wo + (nt * 8 + g) * L::KO + 2 * t;
#pragma unroll
            for (int kb = 0; kb < C / 8; ++kb) mma3(acc, h3[kb], hi + w + 8 * kb, lo + w + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < B / 8; ++kb) mma3(acc, h2[kb], hi + w + C + 8 * kb, lo + w + C + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < A / 8; ++kb) mma3(acc, h1[kb], hi + w + C + B + 8 * kb, lo + w + C + B + 8 * kb);
            if (r0 < rows) __stcs(reinterpret_cast<float2 *>(y + r0 * O + nt * 8 + 2 * t), make_float2(acc[0], acc[1]));
            if (r1 < rows) __stcs(reinterpret_cast<float2 *>(y + r1 * O + nt * 8 + 2 * t), make_float2(acc[2], acc[3]));
        }
    }
}

template <int A, int B, int C, int O>
int launch_fwd(const float *u, const float *w2, const float *w3, const float *wo, int64_t rows, float *y,
               cudaStream_t s) {
    using L = Layout<A, B, C, O>;
    static bool configured = false;  // benign race: the attribute call is idempotent
    if (!configured) {
        PCS_CUDA(cudaFuncSetAttribute(k_dense_mlp3_fwd_mma<A, B, C, O>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)L::bytes));
        configured = true;
    }
    const int64_t n_tiles = (rows + 15) / 16;
    int64_t grid = (n_tiles + kWarps - 1) / kWarps;
    const int64_t cap = 148 * 8;  // grid-stride: the weights are staged (and split) once per CTA
    if (grid > cap) grid = cap;
    k_dense_mlp3_fwd_mma<A, B, C, O><<<(unsigned)grid, kThreads, L::bytes, s>>>(u, w2, w3, wo, rows, y);
    return check_launch("dense_mlp3_forward");
}

// ---- backward ---------------------------------------------------------------------------------------
// Per 16-row tile: recompute h1..h3 (as forward), write hcat = [h3|h2|h1], form dy (dense, or implied by the
// pooled gradient + arg-max as in dense_mlp.cu), then the three dX products against TRANSPOSED weight
// copies (out index = hidden column, in index = the gradient column), masking by h > 0:
//   dpre3 = (dy Wo[:,h3]) * m3;  dpre2 = (dy Wo[:,h2] + dpre3 W3[:,h2]) * m2;
//   dpre1 = (dy Wo[:,h1] + dpre3 W3[:,h1] + dpre2 W2) * m1;          du = [dpre1|dpre2|dpre3|dy].
template <int A, int B, int C, int O>
struct LayoutBwd {
    static constexpr int HC = A + B + C;
    static constexpr int K2 = padk(A), K3 = padk(A + B);            // forward copies (recompute)
    static constexpr int KOT = padk(O), K3T = padk(C), K2T = padk(B);  // transposed copies
    static constexpr int o_w2 = 0, o_w3 = o_w2 + B * K2, o_wot = o_w3 + C * K3, o_w3t = o_wot + HC * KOT,
                         o_w2t = o_w3t + (B + A) * K3T, n_half = o_w2t + A * K2T;
    static constexpr size_t bytes = 2 * sizeof(float) * n_half;
    // in-kernel weight gradients: per-warp staging tile [16][SW] = [dpre3 (C) | dpre2 (B) | h2 (B) | h1 (A)] and
    // the CTA's reduction buffer [NW] = dW3 [C][B+A] then dW2 [B][A]
    static constexpr int SW = padk(C + B + B + A), NW3 = C * (B + A), NW = NW3 + B * A;
    static constexpr size_t bytes_wgrad = bytes + sizeof(float) * (kWarps * 16 * SW + NW);
};

// stage W [N][K] row-major TRANSPOSED: dst row k (stride KP), column n
__device__ __forceinline__ void stage_t(const float *__restrict__ w, int N, int K, int KP, float *hi, float *lo) {
    for (int i = threadIdx.x; i < N * K; i += kThreads) {
        const int n = i / K, k = i - n * K;
        const float v = __ldg(w + i);
        const uint32_t h = to_tf32(v);
        hi[k * KP + n] = __uint_as_float(h);
        lo[k * KP + n] = __uint_as_float(to_tf32(v - __uint_as_float(h)));
    }
}

// WGRAD: also accumulate dW3 = dpre3^T [h2|h1] and dW2 = dpre2^T h1 on the tensor cores (M = weight row,
// N = weight column, K = the tile's 16 rows; operands re-read from a per-warp staging tile in the
// transposed fragment order), one partial per CTA in wpart [gridDim.x][NW].
// CTAs resident per SM: three for the two narrower configurations (168 registers: the kernel waits on loads and on its
// own dependency chains with 8 warps per SM), two for the widest (255 registers)
__host__ __device__ constexpr int bwd_ctas_per_sm(int H) { return H <= 96 ? 3 : 2; }

template <int A, int B, int C, int O, bool POOLED, bool WGRAD>
__global__ void __launch_bounds__(kThreads, bwd_ctas_per_sm(A + B + C + O))
k_dense_mlp3_bwd_mma(const float *__restrict__ u, const float *__restrict__ dy, const int32_t *__restrict__ midx,
                     const int32_t *__restrict__ cidxs, const int32_t *__restrict__ bgs, const float *__restrict__ w2,
                     const float *__restrict__ w3, const float *__restrict__ wo, int64_t rows, float *__restrict__ du,
                     float *__restrict__ hcat, float *__restrict__ wpart) {
    using L = LayoutBwd<A, B, C, O>;
    constexpr int H = A + B + C + O, HC = A + B + C;
    extern __shared__ __align__(16) float smem[];
    float *hi = smem, *lo = smem + L::n_half;
    stage(w2, B, A, L::K2, hi + L::o_w2, lo + L::o_w2);
    stage(w3, C, A + B, L::K3, hi + L::o_w3, lo + L::o_w3);
    stage_t(wo, O, HC, L::KOT, hi + L::o_wot, lo + L::o_wot);
    stage_t(w3, C, A + B, L::K3T, hi + L::o_w3t, lo + L::o_w3t);
    stage_t(w2, B, A, L::K2T, hi + L::o_w2t, lo + L::o_w2t);
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    constexpr int MT3 = (C + 15) / 16, NT3 = (B + A) / 8, MT2 = (B + 15) / 16, NT2 = A / 8;
    float *stg = smem + 2 * L::n_half + warp * 16 * L::SW;  // WGRAD only
    float aw3[WGRAD ? MT3 : 1][WGRAD ? NT3 : 1][4], aw2[WGRAD ? MT2 : 1][WGRAD ? NT2 : 1][4];
    if (WGRAD) {
#pragma unroll
        for (int m = 0; m < MT3; ++m)
#pragma unroll
            for (int n = 0; n < NT3; ++n) aw3[m][n][0] = aw3[m][n][1] = aw3[m][n][2] = aw3[m][n][3] = 0.f;
#pragma unroll
        for (int m = 0; m < MT2; ++m)
#pragma unroll
            for (int n = 0; n < NT2; ++n) aw2[m][n][0] = aw2[m][n][1] = aw2[m][n][2] = aw2[m][n][3] = 0.f;
    }
    // staging: value pair (row g / g+8, columns col+2t, col+2t+1); rows past the end must hold zeros in the d part
#define PCS_STG(col, c0, c1, c2, c3)                                                                   \
    do {                                                                                                  \
        if (WGRAD) {                                                                                      \
            *reinterpret_cast<float2 *>(stg + g * L::SW + (col) + 2 * t) = make_float2(c0, c1);           \
            *reinterpret_cast<float2 *>(stg + (g + 8) * L::SW + (col) + 2 * t) = make_float2(c2, c3);     \
        }                                                                                                 \
    } while (0)
    const int64_t n_tiles = (rows + 15) / 16;
    // Software pipeline (PF, when the register budget allows): the inputs of tile i+1 are in flight while
    // tile i computes, and the centre indices of tile i+2 (which the pooled-gradient addresses of tile
    // i+1 depend on) one step earlier still, so no stage waits on a load it has just issued.
    constexpr bool PF = (H <= 64);
    struct In {
        float2 cu0[HC / 8], cu1[HC / 8];  // pre-activations of rows g / g+8
        float2 gv0[O / 8], gv1[O / 8];    // dy (dense) or the pooled gradient row of the centre
        int2 gm0[POOLED ? O / 8 : 1], gm1[POOLED ? O / 8 : 1];  // arg-max offsets of the centre (POOLED)
        int32_t bg0, bg1;                 // first row of the centre (POOLED)
    };
    const int64_t stride = (int64_t)gridDim.x * kWarps;
    auto clamp_rows = [&](int64_t tile, int64_t &q0, int64_t &q1) {
        const int64_t r0 = tile * 16 + g, r1 = r0 + 8;
        q0 = r0 < rows ? r0 : rows - 1;
        q1 = r1 < rows ? r1 : rows - 1;
    };
    auto load_ids = [&](int64_t tile, int32_t &i0, int32_t &i1) {
        if (!POOLED) return;
        int64_t q0, q1;
        clamp_rows(tile, q0, q1);
        i0 = __ldg(cidxs + q0);
        i1 = __ldg(cidxs + q1);
    };
    auto load_in = [&](int64_t tile, int32_t i0, int32_t i1, In &x) {
        int64_t q0, q1;
        clamp_rows(tile, q0, q1);
        const float *u0 = u + q0 * H + 2 * t;
        const float *u1 = u + q1 * H + 2 * t;
#pragma unroll
        for (int b = 0; b < HC / 8; ++b) {
            x.cu0[b] = __ldg(reinterpret_cast<const float2 *>(u0 + 8 * b));
            x.cu1[b] = __ldg(reinterpret_cast<const float2 *>(u1 + 8 * b));
        }
        if (POOLED) {
            x.bg0 = __ldg(bgs + i0);
            x.bg1 = __ldg(bgs + i1);
#pragma unroll
            for (int b = 0; b < O / 8; ++b) {
                x.gv0[b] = __ldg(reinterpret_cast<const float2 *>(dy + (int64_t)i0 * O + 8 * b + 2 * t));
                x.gv1[b] = __ldg(reinterpret_cast<const float2 *>(dy + (int64_t)i1 * O + 8 * b + 2 * t));
                x.gm0[b] = __ldg(reinterpret_cast<const int2 *>(midx + (int64_t)i0 * O + 8 * b + 2 * t));
                x.gm1[b] = __ldg(reinterpret_cast<const int2 *>(midx + (int64_t)i1 * O + 8 * b + 2 * t));
            }
        } else {
#pragma unroll
            for (int b = 0; b < O / 8; ++b) {
                x.gv0[b] = __ldg(reinterpret_cast<const float2 *>(dy + q0 * O + 8 * b + 2 * t));
                x.gv1[b] = __ldg(reinterpret_cast<const float2 *>(dy + q1 * O + 8 * b + 2 * t));
            }
        }
    };
    In nxt;
    int32_t ni0 = 0, ni1 = 0;  // centre indices of the tile after `nxt`
    const int64_t first = (int64_t)blockIdx.x * kWarps + warp;
    if (PF && first < n_tiles) {
        int32_t a0 = 0, a1 = 0;
        load_ids(first, a0, a1);
        load_in(first, a0, a1, nxt);
        if (first + stride < n_tiles) load_ids(first + stride, ni0, ni1);
    }
    for (int64_t tile = first; tile < n_tiles; tile += stride) {
        const int64_t r0 = tile * 16 + g, r1 = r0 + 8;
        const bool ok0 = r0 < rows, ok1 = r1 < rows;
        const int64_t q0 = ok0 ? r0 : rows - 1, q1 = ok1 ? r1 : rows - 1;
        In cur;
        if (PF) {
            cur = nxt;
            if (tile + stride < n_tiles) {
                load_in(tile + stride, ni0, ni1, nxt);
                if (tile + 2 * stride < n_tiles) load_ids(tile + 2 * stride, ni0, ni1);
            }
        } else {
            int32_t a0 = 0, a1 = 0;
            load_ids(tile, a0, a1);
            load_in(tile, a0, a1, cur);
        }
        float2 (&cu0)[HC / 8] = cur.cu0, (&cu1)[HC / 8] = cur.cu1;
        // dy tile (loads issued before the recompute so that they overlap it)
        float2 gy0[O / 8], gy1[O / 8];
        if (POOLED) {
            const int32_t k0 = (int32_t)(q0 - cur.bg0), k1 = (int32_t)(q1 - cur.bg1);
#pragma unroll
            for (int b = 0; b < O / 8; ++b) {
                gy0[b] = make_float2(cur.gm0[b].x == k0 ? cur.gv0[b].x : 0.f, cur.gm0[b].y == k0 ? cur.gv0[b].y : 0.f);
                gy1[b] = make_float2(cur.gm1[b].x == k1 ? cur.gv1[b].x : 0.f, cur.gm1[b].y == k1 ? cur.gv1[b].y : 0.f);
            }
        } else {
#pragma unroll
            for (int b = 0; b < O / 8; ++b) { gy0[b] = cur.gv0[b]; gy1[b] = cur.gv1[b]; }
        }
        float *hc0 = hcat + r0 * HC + 2 * t, *hc1 = hcat + r1 * HC + 2 * t;
        float *du0 = du + r0 * H + 2 * t, *du1 = du + r1 * H + 2 * t;
        // ---- recompute; pos bit (4*block + i) = (element i of hcat block > 0), blocks in hcat order [h3|h2|h1]
        uint32_t pos = 0;
        Frag h1[A / 8], h2[B / 8];
#define PCS_EMIT_H(blk, c0, c1, c2, c3)                                                          \
        do {                                                                                        \
            if (ok0) __stcs(reinterpret_cast<float2 *>(hc0 + 8 * (blk)), make_float2(c0, c1));      \
            if (ok1) __stcs(reinterpret_cast<float2 *>(hc1 + 8 * (blk)), make_float2(c2, c3));      \
            pos |= ((c0 > 0.f ? 1u : 0u) | (c1 > 0.f ? 2u : 0u) | (c2 > 0.f ? 4u : 0u) | (c3 > 0.f ? 8u : 0u)) << (4 * (blk)); \
        } while (0)
#pragma unroll
        for (int b = 0; b < A / 8; ++b) {
            const float c0 = fmaxf(cu0[b].x, 0.f), c1 = fmaxf(cu0[b].y, 0.f), c2 = fmaxf(cu1[b].x, 0.f), c3 = fmaxf(cu1[b].y, 0.f);
            h1[b] = make_frag(c0, c1, c2, c3);
            PCS_EMIT_H((C + B) / 8 + b, c0, c1, c2, c3);
            PCS_STG(C + B + B + 8 * b, c0, c1, c2, c3);
        }
#pragma unroll
        for (int nt = 0; nt < B / 8; ++nt) {
            float acc[4] = {cu0[A / 8 + nt].x, cu0[A / 8 + nt].y, cu1[A / 8 + nt].x, cu1[A / 8 + nt].y};
            const int w = L::o_w2 + (nt * 8 + g) * L::K2 + 2 * t;
#pragma unroll
            for (int kb = 0; kb < A / 8; ++kb) mma3(acc, h1[kb], hi + w + 8 * kb, lo + w + 8 * kb);
            const float c0 = fmaxf(acc[0], 0.f), c1 = fmaxf(acc[1], 0.f), c2 = fmaxf(acc[2], 0.f), c3 = fmaxf(acc[3], 0.f);
            h2[nt] = make_frag(c0, c1, c2, c3);
            PCS_EMIT_H(C / 8 + nt, c0, c1, c2, c3);
            PCS_STG(C + B + 8 * nt, c0, c1, c2, c3);
        }
#pragma unroll
        for (int nt = 0; nt < C / 8; ++nt) {
            float acc[4] = {cu0[(A + B) / 8 + nt].x, cu0[(A + B) / 8 + nt].y, cu1[(A + B) / 8 + nt].x,
                            cu1[(A + B) / 8 + nt].y};
            const int w = L::o_w3 + (nt * 8 + g) * L::K3 + 2 * t;
#pragma unroll
            for (int kb = 0; kb < B / 8; ++kb) mma3(acc, h2[kb], hi + w + 8 * kb, lo + w + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < A / 8; ++kb) mma3(acc, h1[kb], hi + w + B + 8 * kb, lo + w + B + 8 * kb);
            const float c0 = fmaxf(acc[0], 0.f), c1 = fmaxf(acc[1], 0.f), c2 = fmaxf(acc[2], 0.f), c3 = fmaxf(acc[3], 0.f);
            PCS_EMIT_H(nt, c0, c1, c2, c3);
        }
#undef PCS_EMIT_H
        // ---- dy fragments; d(uo) = dy
        Frag G[O / 8];
#pragma unroll
        for (int b = 0; b < O / 8; ++b) {
            G[b] = make_frag(gy0[b].x, gy0[b].y, gy1[b].x, gy1[b].y);
            if (ok0) __stcs(reinterpret_cast<float2 *>(du0 + HC + 8 * b), gy0[b]);
            if (ok1) __stcs(reinterpret_cast<float2 *>(du1 + HC + 8 * b), gy1[b]);
        }
#define PCS_MASK(blk, acc)                                     \
        do {                                                      \
            const uint32_t mk = pos >> (4 * (blk));               \
            acc[0] = (mk & 1u) ? acc[0] : 0.f; acc[1] = (mk & 2u) ? acc[1] : 0.f; \
            acc[2] = (mk & 4u) ? acc[2] : 0.f; acc[3] = (mk & 8u) ? acc[3] : 0.f; \
        } while (0)
        Frag D3[C / 8], D2[B / 8];
#pragma unroll
        for (int nt = 0; nt < C / 8; ++nt) {  // dpre3: hcat columns [0, C)
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const int w = L::o_wot + (nt * 8 + g) * L::KOT + 2 * t;
#pragma unroll
            for (int kb = 0; kb < O / 8; ++kb) mma3(acc, G[kb], hi + w + 8 * kb, lo + w + 8 * kb);
            PCS_MASK(nt, acc);
            D3[nt] = make_frag(acc[0], acc[1], acc[2], acc[3]);
            PCS_STG(8 * nt, ok0 ? acc[0] : 0.f, ok0 ? acc[1] : 0.f, ok1 ? acc[2] : 0.f, ok1 ? acc[3] : 0.f);
            if (ok0) __stcs(reinterpret_cast<float2 *>(du0 + A + B + 8 * nt), make_float2(acc[0], acc[1]));
            if (ok1) __stcs(reinterpret_cast<float2 *>(du1 + A + B + 8 * nt), make_float2(acc[2], acc[3]));
        }
#pragma unroll
        for (int nt = 0; nt < B / 8; ++nt) {  // dpre2: hcat columns [C, C+B)
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const int wo_ = L::o_wot + (C + nt * 8 + g) * L::KOT + 2 * t;
            const int w3_ = L::o_w3t + (nt * 8 + g) * L::K3T + 2 * t;
#pragma unroll
            for (int kb = 0; kb < O / 8; ++kb) mma3(acc, G[kb], hi + wo_ + 8 * kb, lo + wo_ + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < C / 8; ++kb) mma3(acc, D3[kb], hi + w3_ + 8 * kb, lo + w3_ + 8 * kb);
            PCS_MASK(C / 8 + nt, acc);
            D2[nt] = make_frag(acc[0], acc[1], acc[2], acc[3]);
            PCS_STG(C + 8 * nt, ok0 ? acc[0] : 0.f, ok0 ? acc[1] : 0.f, ok1 ? acc[2] : 0.f, ok1 ? acc[3] : 0.f);
            if (ok0) __stcs(reinterpret_cast<float2 *>(du0 + A + 8 * nt), make_float2(acc[0], acc[1]));
            if (ok1) __stcs(reinterpret_cast<float2 *>(du1 + A + 8 * nt), make_float2(acc[2], acc[3]));
        }
#pragma unroll
        for (int nt = 0; nt < A / 8; ++nt) {  // dpre1: hcat columns [C+B, HC)
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            const int wo_ = L::o_wot + (C + B + nt * 8 + g) * L::KOT + 2 * t;
            const int w3_ = L::o_w3t + (B + nt * 8 + g) * L::K3T + 2 * t;
            const int w2_ = L::o_w2t + (nt * 8 + g) * L::K2T + 2 * t;
#pragma unroll
            for (int kb = 0; kb < O / 8; ++kb) mma3(acc, G[kb], hi + wo_ + 8 * kb, lo + wo_ + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < C / 8; ++kb) mma3(acc, D3[kb], hi + w3_ + 8 * kb, lo + w3_ + 8 * kb);
#pragma unroll
            for (int kb = 0; kb < B / 8; ++kb) mma3(acc, D2[kb], hi + w2_ + 8 * kb, lo + w2_ + 8 * kb);
            PCS_MASK((C + B) / 8 + nt, acc);
            if (ok0) __stcs(reinterpret_cast<float2 *>(du0 + 8 * nt), make_float2(acc[0], acc[1]));
            if (ok1) __stcs(reinterpret_cast<float2 *>(du1 + 8 * nt), make_float2(acc[2], acc[3]));
        }
#undef PCS_MASK
        if (WGRAD) {
            __syncwarp();
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {  // K = rows 8ks .. 8ks+7 of the tile
                const float *sa = stg + (8 * ks + t) * L::SW, *sb = sa + 4 * L::SW;  // rows t and t+4 of this k-step
                // B operands: [h2|h1] columns; b0 = hB[row t][8nt+g], b1 = hB[row t+4][8nt+g]
                uint32_t bh[NT3][2], bl[NT3][2];
#pragma unroll
                for (int nt = 0; nt < NT3; ++nt) {
                    const float v0 = sa[C + B + 8 * nt + g], v1 = sb[C + B + 8 * nt + g];
                    bh[nt][0] = to_tf32(v0); bl[nt][0] = to_tf32(v0 - __uint_as_float(bh[nt][0]));
                    bh[nt][1] = to_tf32(v1); bl[nt][1] = to_tf32(v1 - __uint_as_float(bh[nt][1]));
                }
#pragma unroll
                for (int mt = 0; mt < MT3; ++mt) {  // A = dpre3^T: a0 = d3[row t][16mt+g], a1 = [row t][16mt+g+8], a2/a3 rows t+4
                    const bool up = 16 * mt + g + 8 < C;
                    const Frag a = make_frag(sa[16 * mt + g], sb[16 * mt + g], up ? sa[16 * mt + g + 8] : 0.f,
                                             up ? sb[16 * mt + g + 8] : 0.f);  // make_frag order: (a0, a2, a1, a3)
#pragma unroll
                    for (int nt = 0; nt < NT3; ++nt) {
                        mma_tf32(aw3[mt][nt], a.lo, bh[nt][0], bh[nt][1]);
                        mma_tf32(aw3[mt][nt], a.hi, bl[nt][0], bl[nt][1]);
                        mma_tf32(aw3[mt][nt], a.hi, bh[nt][0], bh[nt][1]);
                    }
                }
#pragma unroll
                for (int mt = 0; mt < MT2; ++mt) {  // A = dpre2^T, B = h1 (the last A/8 column blocks of [h2|h1])
                    const bool up = 16 * mt + g + 8 < B;
                    const Frag a = make_frag(sa[C + 16 * mt + g], sb[C + 16 * mt + g], up ? sa[C + 16 * mt + g + 8] : 0.f,
                                             up ? sb[C + 16 * mt + g + 8] : 0.f);
#pragma unroll
                    for (int nt = 0; nt < NT2; ++nt) {
                        mma_tf32(aw2[mt][nt], a.lo, bh[B / 8 + nt][0], bh[B / 8 + nt][1]);
                        mma_tf32(aw2[mt][nt], a.hi, bl[B / 8 + nt][0], bl[B / 8 + nt][1]);
                        mma_tf32(aw2[mt][nt], a.hi, bh[B / 8 + nt][0], bh[B / 8 + nt][1]);
                    }
                }
            }
            __syncwarp();  // the next tile overwrites the staging rows
        }
    }
#undef PCS_STG
    if (WGRAD) {
        float *red = smem + 2 * L::n_half + kWarps * 16 * L::SW;  // [NW]
        for (int w = 0; w < kWarps; ++w) {  // warp 0 stores, 1..3 add: a fixed order
            if (warp == w) {
#pragma unroll
                for (int mt = 0; mt < MT3; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT3; ++nt)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int m = 16 * mt + g + (e >= 2 ? 8 : 0), n = 8 * nt + 2 * t + (e & 1);
                            if (m < C) { float *q = red + m * (B + A) + n; *q = (w == 0 ? 0.f : *q) + aw3[mt][nt][e]; }
                        }
#pragma unroll
                for (int mt = 0; mt < MT2; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT2; ++nt)
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int m = 16 * mt + g + (e >= 2 ? 8 : 0), n = 8 * nt + 2 * t + (e & 1);
                            if (m < B) { float *q = red + L::NW3 + m * A + n; *q = (w == 0 ? 0.f : *q) + aw2[mt][nt][e]; }
                        }
            }
            __syncthreads();
        }
        for (int i = threadIdx.x; i < L::NW; i += kThreads) wpart[(size_t)blockIdx.x * L::NW + i] = red[i];
    }
}

// two waves of resident CTAs; also the number of dW partials
static inline int64_t bwd_grid(int64_t rows, int H) {
    const int64_t n_tiles = (rows + 15) / 16;
    const int64_t grid = (n_tiles + kWarps - 1) / kWarps;
    const int64_t cap = 148 * 2 * bwd_ctas_per_sm(H);
    return grid > cap ? cap : grid;
}

template <int A, int B, int C, int O>
int launch_bwd(const float *u, const float *dy, const int32_t *midx, const int32_t *cidxs, const int32_t *bgs,
               const float *w2, const float *w3, const float *wo, int64_t rows, float *du, float *hcat, float *wpart,
               cudaStream_t s) {
    using L = LayoutBwd<A, B, C, O>;
    static_assert((A + B + C) / 8 * 4 <= 32, "relu mask is one 32-bit word");
    static bool configured = false;
    if (!configured) {
        PCS_CUDA(cudaFuncSetAttribute(k_dense_mlp3_bwd_mma<A, B, C, O, true, true>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::bytes_wgrad));
        PCS_CUDA(cudaFuncSetAttribute(k_dense_mlp3_bwd_mma<A, B, C, O, true, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::bytes));
        PCS_CUDA(cudaFuncSetAttribute(k_dense_mlp3_bwd_mma<A, B, C, O, false, false>,
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L::bytes));
        configured = true;
    }
    const unsigned grid = (unsigned)bwd_grid(rows, A + B + C + O);
    if (midx && wpart)
        k_dense_mlp3_bwd_mma<A, B, C, O, true, true><<<grid, kThreads, L::bytes_wgrad, s>>>(u, dy, midx, cidxs, bgs, w2, w3,
                                                                                            wo, rows, du, hcat, wpart);
    else if (midx)
        k_dense_mlp3_bwd_mma<A, B, C, O, true, false><<<grid, kThreads, L::bytes, s>>>(u, dy, midx, cidxs, bgs, w2, w3, wo,
                                                                                       rows, du, hcat, nullptr);
    else
        k_dense_mlp3_bwd_mma<A, B, C, O, false, false><<<grid, kThreads, L::bytes, s>>>(u, dy, nullptr, nullptr, nullptr,
                                                                                        w2, w3, wo, rows, du, hcat, nullptr);
    return check_launch("dense_mlp3_backward");
}
}  // namespace

namespace pcs {
// rows > 0, widths already validated by pcs_dense_mlp3_supported
int dense_mlp3_forward_mma(const float *u, const float *w2, const float *w3, const float *wo, int64_t rows, int64_t A,
                           int64_t B, int64_t C, int64_t O, float *y, cudaStream_t s) {
#define PCS_CFG(a, b, c, o) \
    if (A == a && B == b && C == c && O == o) return launch_fwd<a, b, c, o>(u, w2, w3, wo, rows, y, s);
    PCS_CFG(8, 8, 16, 32)
    PCS_CFG(16, 16, 32, 64)
    PCS_CFG(16, 16, 16, 48)
#undef PCS_CFG
    return fail("dense_mlp3: unsupported layer widths (%lld,%lld,%lld,%lld)", (long long)A, (long long)B, (long long)C,
                (long long)O);
}
}  // namespace pcs

namespace pcs {
// dy dense [rows,O] when midx == nullptr, else the pooled gradient [pn,O] with its arg-max (see dense_mlp.cu)
int dense_mlp3_backward_mma(const float *u, const float *dy, const int32_t *midx, const int32_t *cidxs,
                            const int32_t *bgs, const float *w2, const float *w3, const float *wo, int64_t rows,
                            int64_t A, int64_t B, int64_t C, int64_t O, float *du, float *hcat, float *wpart,
                            cudaStream_t s) {
#define PCS_CFG(a, b, c, o) \
    if (A == a && B == b && C == c && O == o)  \
        return launch_bwd<a, b, c, o>(u, dy, midx, cidxs, bgs, w2, w3, wo, rows, du, hcat, wpart, s);
    PCS_CFG(8, 8, 16, 32)
    PCS_CFG(16, 16, 32, 64)
    PCS_CFG(16, 16, 16, 48)
#undef PCS_CFG
    return fail("dense_mlp3: unsupported layer widths (%lld,%lld,%lld,%lld)", (long long)A, (long long)B, (long long)C,
                (long long)O);
}
}  // namespace pcs

namespace pcs {
// number of per-CTA partials the WGRAD backward writes for `rows` rows
int64_t dense_mlp3_wgrad_partials_mma(int64_t rows, int64_t H) { return rows > 0 ? bwd_grid(rows, (int)H) : 0; }
}  // namespace pcs

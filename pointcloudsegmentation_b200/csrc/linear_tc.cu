// linear_tc.cu — point-level linear layers on the 5th-generation tensor cores (tcgen05.mma, accumulators
// in TMEM), fp32 in / fp32 out at fp32 accuracy.
//
//   y[M,N] = x[M,K] · B[N,K]^T (+ bias[N]) (+ addend[M,N])        M ~ 10^5 rows (points), K, N <= ~1500
//
// In the reference these are tf.contrib.layers.fully_connected calls between the custom ops
// (model_pointnet.py:10-24,56-69, model.py:762-799: the per-point halves of pointnet_conv, fc_embed, the
// classifier); stock fp32 GEMMs run them on the FFMA pipe.  Here every product is three TF32 tensor-core
// MMAs on an error-compensated split  a = a_hi + a_lo  (a_hi = tf32(a), a_lo = tf32(a - a_hi)):
//   a·b ~= a_hi·b_hi + a_lo·b_hi + a_hi·b_lo        (the dropped a_lo·b_lo term is ~2^-22 relative)
// accumulated in fp32 in TMEM, so results match an fp32 GEMM to ~1e-6 relative.
//
// One persistent CTA per SM, warp-specialised:
//   warps 0-3   epilogue: tcgen05.ld the 128 x BN accumulator (warp w owns TMEM lanes 32w..32w+31), transpose
//               through shared memory, add bias / addend, store rows coalesced.  The accumulator is double
//               buffered in TMEM (2 x 256 columns), so the epilogue of tile i overlaps the MMAs of tile i+1.
//   warps 4-11  A producers: read the fp32 x tile [128 rows x 32 k] straight from HBM (any K, any row
//               stride: no TMA alignment constraints), split it into hi / lo in registers and store both
//               into the 128-byte-swizzled K-major layout tcgen05 reads; the loads of the next k-block are
//               in flight (registers) while the current one is stored.
//               One producer lane also issues ONE bulk async copy (cp.async.bulk, the TMA engine) per stage
//               for the B operand: the weight is pre-split and pre-swizzled into its shared-memory image by
//               k_linear_tc_prep_b (it is tiny and L2-resident), so a stage of B is one contiguous block.
//   warp 12     one elected lane issues the 12 tcgen05.mma (kind::tf32, M=128, N=BN, K=8) per 32-wide
//               k-block and commits them to the stage's "empty" mbarrier / the tile's "accumulator full" one.
// Pipelines: shared-memory stages full/empty (producers <-> MMA), TMEM accumulators full/empty (MMA <->
// epilogue).  Bound: tensor pipe for N >= 128 (12 x 128 clk of MMA per 16 KB of x), HBM below that.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int BM = 128;             // UMMA M (cta_group::1): accumulator row i lives in TMEM lane i
constexpr int BK = 32;              // fp32 per k-block = 128 bytes = one SWIZZLE_128B row
constexpr int kEpiWarps = 4, kProdWarps = 8;
constexpr int kThreads = (kEpiWarps + kProdWarps + 1) * 32;  // 416
constexpr int kMaxStages = 6;
constexpr int kATile = BM * BK * 4;  // 16384 bytes (one of hi / lo)
constexpr int kStgPitch = 33;
constexpr int kStgBytes = kEpiWarps * 32 * kStgPitch * 4;  // 16896
constexpr int kBarBytes = 256;
constexpr int kSmemMax = 232448;  // 227 KB
constexpr uint32_t kTmemCols = 512;
constexpr int kAccCols = 256;  // TMEM columns per accumulator stage

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t to_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }

// ---- mbarrier ------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// A lost arrival would otherwise hang the GPU: give up (trap -> launch error) after ~seconds of polling.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try(bar, parity)) {
        if (++spins > (1u << 24)) __trap();
    }
}

// ---- tcgen05 ---------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// all prior tcgen05.mma of this thread -> one arrival on `bar` when they have completed
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// K-major, SWIZZLE_128B operand tile: rows of 128 bytes, 8-row atoms 1024 bytes apart (SBO), LBO unused (=1),
// descriptor version 1 (sm_100), layout type 2 (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

struct LinParams {
    const float *x;
    int64_t ldx;
    int M, K;
    const float *img;  // [NT][KB][2][BN][32] swizzled hi / lo image of B
    int N, BN, NT, KB;
    const float *bias;
    const float *add;
    int64_t ldadd;
    float *y;
    int64_t ldy;
    int act;  // 0 none, 1 relu
    int stages, num_tiles;
    int dbg;  // development switches (tools/ab_linear_tc.py): 1 no B copy, 2 one MMA per product, 4 no stores, 8 no x loads
};

// ---- weight image ----------------------------------------------------------------------------------
// img[nt][kb][h][n][32 floats]: exactly the bytes a shared-memory stage of B holds (128-byte rows, 16-byte
// chunks XOR-swizzled by row % 8), zero padded in n and k, h = 0 hi / 1 lo.  B[n,k] = w[n*sn + k*sk].
__global__ void __launch_bounds__(256)
k_linear_tc_prep_b(const float *__restrict__ w, int64_t sn, int64_t sk, int N, int K, int BN, int NT, int KB,
                   float *__restrict__ img) {
    const int64_t total = (int64_t)NT * KB * BN * BK;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int k_in = (int)(idx % BK);
        int64_t t = idx / BK;
        const int n_in = (int)(t % BN);
        t /= BN;
        const int kb = (int)(t % KB);
        const int nt = (int)(t / KB);
        const int n = nt * BN + n_in, k = kb * BK + k_in;
        const float v = (n < N && k < K) ? __ldg(w + (int64_t)n * sn + (int64_t)k * sk) : 0.f;
        const uint32_t hi = to_tf32(v);
        const uint32_t lo = to_tf32(v - __uint_as_float(hi));
        const int off = n_in * BK + ((((k_in >> 2) ^ (n_in & 7)) << 2) | (k_in & 3));
        float *base = img + ((int64_t)(nt * KB + kb) * 2) * BN * BK;
        base[off] = __uint_as_float(hi);
        base[BN * BK + off] = __uint_as_float(lo);
    }
}

// ---- main kernel -----------------------------------------------------------------------------------
template <bool VEC>
__global__ void __launch_bounds__(kThreads, 1)
k_linear_tc(const LinParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);  // SWIZZLE_128B atoms need 1024-byte alignment
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int stage_bytes = 2 * kATile + p.BN * 256;
    float *stg = reinterpret_cast<float *>(sm + p.stages * stage_bytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(stg) + kStgBytes);
    const uint32_t bar0 = smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
    auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + a); };
    auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + 2 + a); };
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 4);

    if (threadIdx.x == kThreads - 32) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(full_bar(s), kProdWarps + 1);  // one arrival per producer warp + the expect_tx arrival for B
            mbar_init(empty_bar(s), 1);              // tcgen05.commit
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(tfull_bar(a), 1);
            mbar_init(tempty_bar(a), kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(kTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp >= kEpiWarps && warp < kEpiWarps + kProdWarps) {
        // ================= A producers (+ the B bulk copy) =================
        const int pw = warp - kEpiWarps;
        float4 cur[4], nxt[4];
        auto load = [&](float4 (&r)[4], int tile, int kb) {
            const int mt = tile / p.NT;
            if (VEC) {  // lane -> (row lane/8 of 4, 16-byte chunk lane%8)
                const int col = kb * BK + (lane & 7) * 4;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int row = mt * BM + pw * 16 + i * 4 + (lane >> 3);
                    r[i] = (row < p.M && col < p.K) ? __ldg(reinterpret_cast<const float4 *>(p.x + (int64_t)row * p.ldx + col))
                                                    : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            } else {  // lane -> column, 16 rows per warp
                const int col = kb * BK + lane;
                const bool cok = col < p.K;
                float *f = reinterpret_cast<float *>(r);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int row = mt * BM + pw * 16 + i;
                    f[i] = (cok && row < p.M) ? __ldg(p.x + (int64_t)row * p.ldx + col) : 0.f;
                }
            }
        };
        int tile = blockIdx.x, kb = 0;
        uint32_t it = 0;
        if (tile < p.num_tiles) load(cur, tile, 0);
        while (tile < p.num_tiles) {
            int ntile = tile, nkb = kb + 1;
            if (nkb == p.KB) { nkb = 0; ntile += gridDim.x; }
            if (ntile < p.num_tiles && !(p.dbg & 8)) load(nxt, ntile, nkb);  // in flight while this k-block is split and stored
            const int s = (int)(it % (uint32_t)p.stages);
            const uint32_t ph = (it / (uint32_t)p.stages) & 1u;
            mbar_wait(empty_bar(s), ph ^ 1u);
            uint8_t *a_hi = sm + s * stage_bytes, *a_lo = a_hi + kATile;
            if (pw == 0 && lane == 0) {
                const uint32_t bytes = (uint32_t)p.BN * 256u;
                const int nt = tile % p.NT;
                const float *src = p.img + ((int64_t)(nt * p.KB + kb) * 2) * p.BN * BK;
                if (p.dbg & 1) mbar_arrive(full_bar(s));
                else {
                mbar_arrive_expect_tx(full_bar(s), bytes);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(a_hi + 2 * kATile)), "l"(src), "r"(bytes), "r"(full_bar(s)) : "memory");
                }
            }
            if (VEC) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = pw * 16 + i * 4 + (lane >> 3);
                    const int off = r * 128 + (((lane & 7) ^ (r & 7)) << 4);
                    const float4 v = cur[i];
                    uint4 h, l;
                    h.x = to_tf32(v.x); h.y = to_tf32(v.y); h.z = to_tf32(v.z); h.w = to_tf32(v.w);
                    l.x = to_tf32(v.x - __uint_as_float(h.x)); l.y = to_tf32(v.y - __uint_as_float(h.y));
                    l.z = to_tf32(v.z - __uint_as_float(h.z)); l.w = to_tf32(v.w - __uint_as_float(h.w));
                    *reinterpret_cast<uint4 *>(a_hi + off) = h;
                    *reinterpret_cast<uint4 *>(a_lo + off) = l;
                }
            } else {
                const float *f = reinterpret_cast<const float *>(cur);
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int r = pw * 16 + i;
                    const int off = r * 128 + ((((lane >> 2) ^ (r & 7)) << 4) | ((lane & 3) << 2));
                    const uint32_t h = to_tf32(f[i]);
                    *reinterpret_cast<uint32_t *>(a_hi + off) = h;
                    *reinterpret_cast<uint32_t *>(a_lo + off) = to_tf32(f[i] - __uint_as_float(h));
                }
            }
            fence_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
            __syncwarp();
            if (lane == 0) mbar_arrive(full_bar(s));
#pragma unroll
            for (int i = 0; i < 4; ++i) cur[i] = nxt[i];
            tile = ntile; kb = nkb; ++it;
        }
    } else if (warp == kEpiWarps + kProdWarps) {
        // ================= MMA issuer: one thread =================
        if (lane == 0) {
            // cute::UMMA::InstrDescriptor: c_format F32 (1<<4), a/b_format TF32 (2<<7, 2<<10), K-major both, N>>3, M>>4
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            uint32_t it = 0, tl = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++tl) {
                const int a = (int)(tl & 1u);
                mbar_wait(tempty_bar(a), ((tl >> 1) & 1u) ^ 1u);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d = tmem_base + (uint32_t)(a * kAccCols);
                for (int kb = 0; kb < p.KB; ++kb, ++it) {
                    const int s = (int)(it % (uint32_t)p.stages);
                    mbar_wait(full_bar(s), (it / (uint32_t)p.stages) & 1u);
                    tc_fence_after();
                    const uint32_t a_hi = smem_u32(sm + s * stage_bytes);
                    const uint64_t dah = umma_desc(a_hi), dal = umma_desc(a_hi + kATile);
                    const uint64_t dbh = umma_desc(a_hi + 2 * kATile), dbl = umma_desc(a_hi + 2 * kATile + p.BN * 128);
#pragma unroll
                    for (int k = 0; k < BK / 8; ++k) {  // UMMA K = 8 tf32 = 32 bytes: +2 in the (addr >> 4) field
                        if (p.dbg & 2) { tc_mma_tf32(d, dah + 2 * k, dbh + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u); continue; }
                        tc_mma_tf32(d, dal + 2 * k, dbh + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                        tc_mma_tf32(d, dah + 2 * k, dbl + 2 * k, idesc, 1u);
                        tc_mma_tf32(d, dah + 2 * k, dbh + 2 * k, idesc, 1u);
                    }
                    tc_commit(empty_bar(s));  // stage reusable once these MMAs have read it
                }
                tc_commit(tfull_bar(a));  // accumulator complete
            }
        }
    } else {
        // ================= epilogue: TMEM -> registers -> smem transpose -> coalesced rows =================
        float *my = stg + warp * 32 * kStgPitch;
        uint32_t tl = 0;
        for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++tl) {
            const int mt = tile / p.NT, nt = tile % p.NT;
            const int a = (int)(tl & 1u);
            mbar_wait(tfull_bar(a), (tl >> 1) & 1u);
            tc_fence_after();
            const int row_base = mt * BM + warp * 32;
            for (int c = 0; c < p.BN / 32; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(a * kAccCols + c * 32), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) my[lane * kStgPitch + j] = __uint_as_float(v[j]);
                __syncwarp();
                const int col = nt * p.BN + c * 32 + lane;
                if (col < p.N && !(p.dbg & 4)) {
                    const float b = p.bias ? __ldg(p.bias + col) : 0.f;
#pragma unroll 8
                    for (int r = 0; r < 32; ++r) {
                        const int row = row_base + r;
                        if (row < p.M) {
                            float o = my[r * kStgPitch + lane] + b;
                            if (p.add) o += __ldg(p.add + (int64_t)row * p.ldadd + col);
                            if (p.act == 1) o = fmaxf(o, 0.f);
                            p.y[(int64_t)row * p.ldy + col] = o;
                        }
                    }
                }
                __syncwarp();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(a));
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

struct Tiling {
    int NT, BN, KB, stages;
    size_t img_bytes, smem_bytes;
};
Tiling make_tiling_tc(int64_t N, int64_t K) {
    Tiling t;
    t.NT = (int)((N + 255) / 256);
    const int64_t per = (N + t.NT - 1) / t.NT;
    t.BN = (int)((per + 31) / 32 * 32);
    t.KB = (int)((K + BK - 1) / BK);
    const int stage_bytes = 2 * kATile + t.BN * 256;
    const int avail = kSmemMax - kStgBytes - kBarBytes - 1024;
    t.stages = avail / stage_bytes;
    if (t.stages > kMaxStages) t.stages = kMaxStages;
    t.smem_bytes = (size_t)t.stages * stage_bytes + kStgBytes + kBarBytes + 1024;
    t.img_bytes = (size_t)t.NT * t.KB * 2 * t.BN * BK * 4;
    return t;
}
}  // namespace

extern "C" size_t pcs_linear_tc_workspace_bytes(int64_t N, int64_t K) {
    if (N <= 0 || K <= 0) return 0;
    return make_tiling_tc(N, K).img_bytes;
}

extern "C" int pcs_linear_tc(const float *x, int64_t M, int64_t K, int64_t ldx, const float *w, int64_t N,
                             int64_t w_stride_n, int64_t w_stride_k, const float *bias, const float *addend,
                             int64_t ldadd, int act, float *y, int64_t ldy, void *workspace, size_t workspace_bytes,
                             pcs_stream_t stream) {
    PCS_REQUIRE(M >= 0 && K > 0 && N > 0, "linear_tc: bad sizes M=%lld K=%lld N=%lld", (long long)M, (long long)K, (long long)N);
    PCS_REQUIRE(M < (1ll << 31) - BM && K < (1 << 20) && N < (1 << 20), "linear_tc: sizes out of range");
    PCS_REQUIRE(ldx >= K && ldy >= N && (!addend || ldadd >= N), "linear_tc: row strides smaller than the row");
    const int dbg = act >> 8;
    act &= 0xff;
    PCS_REQUIRE(act == 0 || act == 1, "linear_tc: act must be 0 (none) or 1 (relu)");
    if (M == 0) return 0;
    PCS_REQUIRE(x && w && y, "linear_tc: null pointer");
    const Tiling t = make_tiling_tc(N, K);
    PCS_REQUIRE(workspace && workspace_bytes >= t.img_bytes && aligned(workspace, 16),
                "linear_tc: workspace must be 16-byte aligned and >= %zu bytes", t.img_bytes);
    cudaStream_t s = as_stream(stream);
    float *img = static_cast<float *>(workspace);
    {
        const int64_t total = (int64_t)t.NT * t.KB * t.BN * BK;
        const int grid = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
        k_linear_tc_prep_b<<<grid, 256, 0, s>>>(w, w_stride_n, w_stride_k, (int)N, (int)K, t.BN, t.NT, t.KB, img);
        if (int e = check_launch("linear_tc prep")) return e;
    }
    LinParams p;
    p.x = x; p.ldx = ldx; p.M = (int)M; p.K = (int)K;
    p.img = img; p.N = (int)N; p.BN = t.BN; p.NT = t.NT; p.KB = t.KB;
    p.bias = bias; p.add = addend; p.ldadd = ldadd; p.y = y; p.ldy = ldy; p.act = act;
    p.stages = t.stages;
    p.dbg = dbg;
    const int64_t tiles = ((M + BM - 1) / BM) * t.NT;
    PCS_REQUIRE(tiles < (1ll << 31), "linear_tc: too many tiles");
    p.num_tiles = (int)tiles;
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        PCS_CUDA(cudaGetDevice(&dev));
        PCS_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const int grid = (int)(tiles < sm_count ? tiles : sm_count);
    const bool vec = (K % 4 == 0) && (ldx % 4 == 0) && aligned(x, 16);
    if (vec) {
        PCS_CUDA(cudaFuncSetAttribute(k_linear_tc<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smem_bytes));
        k_linear_tc<true><<<grid, kThreads, t.smem_bytes, s>>>(p);
    } else {
        PCS_CUDA(cudaFuncSetAttribute(k_linear_tc<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smem_bytes));
        k_linear_tc<false><<<grid, kThreads, t.smem_bytes, s>>>(p);
    }
    return check_launch("linear_tc");
}

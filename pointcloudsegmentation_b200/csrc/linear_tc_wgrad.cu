// linear_tc_wgrad.cu — weight (and bias) gradient of the point-level linear layers on the tcgen05 tensor
// cores, fp32 in / fp32 out at fp32 accuracy (the companion of linear_tc.cu).
//
//   dW[N,K] = g[M,N]^T · x[M,K]        (+ db[N] = column sums of g, as one extra "ones" column of x)
//
// a reduction over the M ~ 10^5 - 5·10^5 point rows.  In the reference this is the gradient TensorFlow derives
// for fully_connected (model_pointnet.py:10-24,56-69, model.py:762-799); stock fp32 GEMMs put a handful of
// 64x64 output tiles on 148 SMs for it, or need the reduction cut into slabs (fused_ops.tn_tall).
//
// Both operands are read in their natural row-major form: for this product the contraction index is the ROW,
// so g^T and x^T are "MN-major" UMMA operands (the 128-byte swizzle atom is 8 rows x 32 consecutive columns)
// and a tile row goes to shared memory exactly as it lies in HBM — no transpose anywhere.  As in linear_tc.cu
// every product is three TF32 MMAs on an error-compensated hi / lo split, accumulated in fp32 in TMEM.
//
// One CTA per (128 output rows n) x (BK <= 256 output columns k) x (row split s): it walks its rows in stages of
// 16 (two UMMA k-steps of 8 rows), 4-6 stages deep:
//   warps 4-19  producers: each warp reads 1 row of g (128 columns) and of x (BK columns) per stage from HBM,
//               splits hi / lo in registers and stores both into the swizzled MN-major tiles; loads of the next
//               THREE stages are in flight in registers; all stage-invariant address work is hoisted.
//   warp 0      first walks the pipeline as the MMA issuer — one elected lane issues 6 tcgen05.mma (kind::tf32, M=128,
//               N=BK, K=8) per stage and commits them to the stage's "empty" mbarrier, after the last stage to the
//               "accumulator full" one — then joins the epilogue.  (20 warps: a separate issuer warp made it 21, which
//               cost the registers of 24 — 80 per thread instead of 96 — and the producers are the critical role:
//               tools/wgrad_probe.py, the tensor pipe is busy 38 % of the launch on the largest shape of a step.)
//   warps 0-3   epilogue (once per CTA): tcgen05.ld the 128 x BK accumulator -> smem transpose -> partial[s][n][k]
//               rows stored coalesced.
// The S partials are folded by pcs_column_sum (fixed order: deterministic, no atomics).
#include "common.cuh"
#include "tcgen05.cuh"

using namespace pcs;
using namespace pcs::tc;

namespace {
constexpr int TN = 128;      // output rows per CTA = UMMA M
constexpr int RS = 16;       // contraction rows per stage (two UMMA k-steps)
constexpr int kEpiWarps = 4, kProdWarps = 16;  // one producer warp per row of a stage
// 20 warps: the MMA issuer is epilogue warp 0 (idle until the accumulator is complete).  A separate 21st warp made the CTA
// cost the registers of 24 (80 per thread); with 20 the producers have 96 and keep a fourth load set in flight.
constexpr int kThreads = (kEpiWarps + kProdWarps) * 32;  // 640
constexpr int kMaxStages = 6;
constexpr int kATile = RS * TN * 4;  // 8192 bytes (one of hi / lo)
constexpr int kStgPitch = 20;
constexpr int kStgBytes = kEpiWarps * 32 * kStgPitch * 4;  // 10240
constexpr int kBarBytes = 256;
constexpr int kSmemMax = 232448;
constexpr uint32_t kTmemCols = 256;

// MN-major TF32 operands have ONE legal shared-memory layout, SWIZZLE_128B_BASE32B (cutlass sm100_common.inl:92;
// with plain SWIZZLE_128B the MMA silently yields zeros): atoms of 4 rows x 128 bytes (32 columns), 32-byte
// pieces XOR-swizzled by the row (cute Swizzle<2,5,2>); atoms along the columns LBO = 512 bytes apart, along
// the rows (4 at a time) SBO = W/32 * 512 bytes apart.  A UMMA k-step (8 rows) spans two row atoms.
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(512u >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) |
           (1ull << 46) | (1ull << 61);
}
// byte offset of (row mi of the stage, column e) in a tile W columns wide
__device__ __forceinline__ int mn_off(int mi, int e, int W) {
    const int r = mi & 3;
    return (mi >> 2) * (W >> 5) * 512 + (e >> 5) * 512 + r * 128 + (((((e & 31) >> 3) ^ r)) << 5) + ((e & 7) << 2);
}

struct WgParams {
    const float *g;
    int64_t ldg;
    const float *x;
    int64_t ldx;
    int M, N, K, Kext;  // Kext = K + 1 with the ones column (bias gradient), else K
    int BK, KT, NT, S;  // output column tile, # column tiles, # row tiles (of 128), # row splits
    int rows_per_split;  // multiple of RS
    float *partial;      // [S][N][Kext]
    int stages;
    int dbg;  // development: 1 = issue every UMMA three times (wrong results) to measure its cost
};

template <bool AG, bool AX>  // g / x row pieces 16-byte aligned
__global__ void __launch_bounds__(kThreads, 1)
k_linear_tc_wgrad(const WgParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b_tile = RS * p.BK * 4;  // bytes of one of hi / lo of the x tile
    const int stage_bytes = 2 * kATile + 2 * b_tile;
    float *stg = reinterpret_cast<float *>(sm + p.stages * stage_bytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(stg) + kStgBytes);
    const uint32_t bar0 = smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
    const uint32_t tfull_bar = bar0 + 8u * (2 * kMaxStages);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 1);

    int id = blockIdx.x;
    const int kt = id % p.KT; id /= p.KT;
    const int nt = id % p.NT;
    const int sp = id / p.NT;
    const int n0 = nt * TN, k0 = kt * p.BK;
    const int m_beg = sp * p.rows_per_split;
    const int m_end = min(p.M, m_beg + p.rows_per_split);
    const int n_stages = (m_end - m_beg + RS - 1) / RS;  // >= 1 by construction

    if (threadIdx.x == kThreads - 32) {
        for (int s = 0; s < p.stages; ++s) {
            mbar_init(full_bar(s), kProdWarps);
            mbar_init(empty_bar(s), 1);
        }
        mbar_init(tfull_bar, 1);
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
        // ================= producers: warp pw owns row pw of every 16-row stage =================
        // ncu: the producers were bound by their own instruction stream (~300 dependent instructions per warp and
        // stage), not by memory: so one row per warp, and everything that does not change from stage to stage —
        // swizzled shared-memory offsets, column predicates, the lane's global pointers — is computed once.
        const int pw = warp - kEpiWarps;  // = row of the stage
        constexpr int kX = 2;             // float4 registers for x per stage (BK <= 256)
        // four register sets in rotation, indexed by compile-time constants (loop unrolled by 4): loads run THREE stages
        // ahead (two stages ahead the kernel still waited on them; one stage ahead costs 10-18 %)
        float4 RG[4], RX[4][kX];
        // lane -> g columns 4*lane .. +3, x columns q*128 + 4*lane .. +3 (q < 2): one 16-byte load when the row
        // pieces are 16-byte aligned (AG / AX), four 4-byte loads otherwise (odd K such as 643, 131, 35: the
        // stores stay 16 bytes wide either way)
        const int offA = mn_off(pw, lane * 4, TN);
        int offB[kX];
        uint32_t okB = 0, tailB = 0;  // per-register: column inside the tile / piece holds the ones column or padding
        const int gcols = min(4, p.N - (n0 + lane * 4));  // valid g columns of this lane's piece (<= 0: none)
#pragma unroll
        for (int q = 0; q < kX; ++q) {
            const int e = q * 128 + lane * 4;
            offB[q] = e < p.BK ? mn_off(pw, e, p.BK) : 0;
            if (e < p.BK) {
                okB |= 1u << q;
                if (k0 + e + 3 >= p.K) tailB |= 1u << q;
            }
        }
        const float *gp = p.g + (int64_t)(m_beg + pw) * p.ldg + n0 + lane * 4;
        const float *xp = p.x + (int64_t)(m_beg + pw) * p.ldx + k0 + lane * 4;
        const int64_t gstep = (int64_t)RS * p.ldg, xstep = (int64_t)RS * p.ldx;
        auto load = [&](float4 &rg, float4 (&rx)[kX], int st) {
            const bool mok = m_beg + st * RS + pw < m_end;
            const float *g = gp + (int64_t)st * gstep, *x = xp + (int64_t)st * xstep;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (mok && gcols > 0) {
                if (AG && gcols == 4) v = __ldg(reinterpret_cast<const float4 *>(g));
                else {
                    v.x = __ldg(g);
                    if (gcols > 1) v.y = __ldg(g + 1);
                    if (gcols > 2) v.z = __ldg(g + 2);
                    if (gcols > 3) v.w = __ldg(g + 3);
                }
            }
            rg = v;
#pragma unroll
            for (int q = 0; q < kX; ++q) {
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (mok && ((okB >> q) & 1u)) {
                    if (!((tailB >> q) & 1u)) {
                        if (AX) w = __ldg(reinterpret_cast<const float4 *>(x + q * 128));
                        else w = make_float4(__ldg(x + q * 128), __ldg(x + q * 128 + 1), __ldg(x + q * 128 + 2), __ldg(x + q * 128 + 3));
                    } else {
                        const int k = k0 + q * 128 + lane * 4;
                        float t[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) t[u] = (k + u < p.K) ? __ldg(x + q * 128 + u) : ((k + u < p.Kext) ? 1.f : 0.f);
                        w = make_float4(t[0], t[1], t[2], t[3]);
                    }
                }
                rx[q] = w;
            }
        };
        auto split_store4 = [&](uint8_t *hi, uint8_t *lo, int off, const float4 v) {
            uint4 hh, ll;
            hh.x = to_tf32(v.x); hh.y = to_tf32(v.y); hh.z = to_tf32(v.z); hh.w = to_tf32(v.w);
            ll.x = __float_as_uint(v.x - __uint_as_float(hh.x)); ll.y = __float_as_uint(v.y - __uint_as_float(hh.y));
            ll.z = __float_as_uint(v.z - __uint_as_float(hh.z)); ll.w = __float_as_uint(v.w - __uint_as_float(hh.w));
            *reinterpret_cast<uint4 *>(hi + off) = hh;
            *reinterpret_cast<uint4 *>(lo + off) = ll;
        };
        load(RG[0], RX[0], 0);
        if (n_stages > 1) load(RG[1], RX[1], 1);
        if (n_stages > 2) load(RG[2], RX[2], 2);
        uint32_t s = 0, ph = 0;
        for (int st0 = 0; st0 < n_stages; st0 += 4) {
#pragma unroll
            for (int q3 = 0; q3 < 4; ++q3) {
                const int st = st0 + q3;
                if (st >= n_stages) break;
                if (st + 3 < n_stages) load(RG[(q3 + 3) % 4], RX[(q3 + 3) % 4], st + 3);  // three stages ahead in registers
                mbar_wait<20>(empty_bar((int)s), ph ^ 1u);
                uint8_t *a_hi = sm + s * stage_bytes, *a_lo = a_hi + kATile;
                uint8_t *b_hi = a_hi + 2 * kATile, *b_lo = b_hi + b_tile;
                split_store4(a_hi, a_lo, offA, RG[q3]);
#pragma unroll
                for (int q = 0; q < kX; ++q)
                    if ((okB >> q) & 1u) split_store4(b_hi, b_lo, offB[q], RX[q3][q]);
                fence_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(full_bar((int)s));
                if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
            }
        }
    } else {
      if (warp == 0) {
        // ================= MMA issuer (then this warp joins the epilogue) =================
        // InstrDescriptor: c F32, a/b TF32, a_major = b_major = MN (bits 15, 16), N >> 3, M >> 4
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (1u << 15) | (1u << 16) |
                               ((uint32_t)(p.BK >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t a_sbo = (TN >> 5) * 512, b_sbo = (uint32_t)(p.BK >> 5) * 512;  // 4-row atom stride
        const uint32_t a_kstep = 2 * a_sbo, b_kstep = 2 * b_sbo;                      // 8 rows per UMMA k-step
        // lean loop (see linear_tc.cu): descriptors of stage 0 built once, + s * stage_bytes / 16 per stage
        const uint32_t base0 = smem_u32(sm);
        const uint64_t dah0 = umma_desc_mn(base0, a_sbo), dal0 = umma_desc_mn(base0 + kATile, a_sbo);
        const uint64_t dbh0 = umma_desc_mn(base0 + 2 * kATile, b_sbo), dbl0 = umma_desc_mn(base0 + 2 * kATile + b_tile, b_sbo);
        const uint64_t dstep = (uint64_t)(stage_bytes >> 4), ak = (uint64_t)(a_kstep >> 4), bk = (uint64_t)(b_kstep >> 4);
        uint32_t s = 0, ph = 0;
        uint64_t soff = 0;
        for (int st = 0; st < n_stages; ++st) {
            mbar_wait<0>(full_bar((int)s), ph);
            tc_fence_after();
            if (elect_one()) {
                const uint64_t dah = dah0 + soff, dal = dal0 + soff, dbh = dbh0 + soff, dbl = dbl0 + soff;
                tc_mma_tf32(tmem_base, dal, dbh, idesc, st > 0 ? 1u : 0u);
                tc_mma_tf32(tmem_base, dah, dbl, idesc, 1u);
                tc_mma_tf32(tmem_base, dah, dbh, idesc, 1u);
                tc_mma_tf32(tmem_base, dal + ak, dbh + bk, idesc, 1u);
                tc_mma_tf32(tmem_base, dah + ak, dbl + bk, idesc, 1u);
                tc_mma_tf32(tmem_base, dah + ak, dbh + bk, idesc, 1u);
                if (p.dbg & 1) {
                    for (int rep = 0; rep < 2; ++rep) {
                        tc_mma_tf32(tmem_base, dal, dbh, idesc, 1u); tc_mma_tf32(tmem_base, dah, dbl, idesc, 1u);
                        tc_mma_tf32(tmem_base, dah, dbh, idesc, 1u); tc_mma_tf32(tmem_base, dal + ak, dbh + bk, idesc, 1u);
                        tc_mma_tf32(tmem_base, dah + ak, dbl + bk, idesc, 1u); tc_mma_tf32(tmem_base, dah + ak, dbh + bk, idesc, 1u);
                    }
                }
                tc_commit(empty_bar((int)s));
                if (st == n_stages - 1) tc_commit(tfull_bar);
            }
            __syncwarp();
            if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; soff = 0; } else soff += dstep;
        }
      }
      {
        // ================= epilogue: the partial tile of this split =================
        float *my = stg + warp * 32 * kStgPitch;
        mbar_wait<200>(tfull_bar, 0u);
        tc_fence_after();
        float *out = p.partial + (int64_t)sp * p.N * p.Kext;
        for (int c = 0; c < p.BK / 16; ++c) {
            uint32_t v[16];
            tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(c * 16), v);
#pragma unroll
            for (int j = 0; j < 4; ++j)
                *reinterpret_cast<uint4 *>(my + lane * kStgPitch + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            __syncwarp();
            const int k = k0 + c * 16 + (lane & 15);
            float o[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) o[i] = my[(i * 2 + (lane >> 4)) * kStgPitch + (lane & 15)];
            if (k < p.Kext) {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int n = n0 + warp * 32 + i * 2 + (lane >> 4);
                    if (n < p.N) out[(int64_t)n * p.Kext + k] = o[i];
                }
            }
            __syncwarp();
        }
      }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

struct WgTiling {
    int Kext, BK, KT, NT, S, rows_per_split, stages;
    size_t smem_bytes;
};
WgTiling make_wg_tiling(int64_t M, int64_t N, int64_t K, int with_bias, int sm_count) {
    WgTiling t;
    t.Kext = (int)K + (with_bias ? 1 : 0);
    t.KT = (t.Kext + 255) / 256;
    const int per = (t.Kext + t.KT - 1) / t.KT;
    t.BK = (per + 31) / 32 * 32;  // whole 32-column swizzle atoms
    t.NT = (int)((N + TN - 1) / TN);
    const int tiles = t.KT * t.NT;
    // row splits: fill the SMs (one CTA each), at least 4 stages of work per CTA
    int64_t S = (2 * sm_count + tiles - 1) / tiles;  // ~2 CTAs per SM over the launch evens out the tail
    const int64_t max_s = (M + 4 * RS - 1) / (4 * RS);
    if (S > max_s) S = max_s;
    if (S < 1) S = 1;
    int64_t rps = (M + S - 1) / S;
    rps = (rps + RS - 1) / RS * RS;
    t.rows_per_split = (int)rps;
    t.S = (int)((M + rps - 1) / rps);
    const int stage_bytes = 2 * kATile + 2 * RS * t.BK * 4;
    const int avail = kSmemMax - kStgBytes - kBarBytes - 1024;
    t.stages = avail / stage_bytes;
    if (t.stages > kMaxStages) t.stages = kMaxStages;
    t.smem_bytes = (size_t)t.stages * stage_bytes + kStgBytes + kBarBytes + 1024;
    return t;
}
int sm_count_cached() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) n = 148;
    }
    return n;
}
}  // namespace

// number of row splits S: `partial` must hold S*N*(K + with_bias) floats; fold them with pcs_column_sum
extern "C" int64_t pcs_linear_tc_wgrad_partials(int64_t M, int64_t N, int64_t K, int with_bias) {
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    return make_wg_tiling(M, N, K, with_bias, sm_count_cached()).S;
}

extern "C" int pcs_linear_tc_wgrad(const float *g, int64_t M, int64_t N, int64_t ldg, const float *x, int64_t K, int64_t ldx,
                                   int with_bias, float *partial, pcs_stream_t stream) {
    const int dbg = with_bias >> 8;
    with_bias &= 1;
    PCS_REQUIRE(M > 0 && N > 0 && K > 0, "linear_tc_wgrad: bad sizes M=%lld N=%lld K=%lld", (long long)M, (long long)N, (long long)K);
    PCS_REQUIRE(M < (1ll << 31) - 4096 && K < (1 << 20) && N < (1 << 20), "linear_tc_wgrad: sizes out of range");
    PCS_REQUIRE(ldg >= N && ldx >= K, "linear_tc_wgrad: row strides smaller than the row");
    PCS_REQUIRE(g && x && partial, "linear_tc_wgrad: null pointer");
    const WgTiling t = make_wg_tiling(M, N, K, with_bias, sm_count_cached());
    WgParams p;
    p.g = g; p.ldg = ldg; p.x = x; p.ldx = ldx;
    p.M = (int)M; p.N = (int)N; p.K = (int)K; p.Kext = t.Kext;
    p.BK = t.BK; p.KT = t.KT; p.NT = t.NT; p.S = t.S; p.rows_per_split = t.rows_per_split;
    p.partial = partial; p.stages = t.stages; p.dbg = dbg;
    const int64_t grid = (int64_t)t.KT * t.NT * t.S;
    PCS_REQUIRE(grid < (1ll << 31), "linear_tc_wgrad: grid too large");
    cudaStream_t s = as_stream(stream);
    const bool ag = (ldg % 4 == 0) && aligned(g, 16), ax = (ldx % 4 == 0) && aligned(x, 16);
#define PCS_WG(AG_, AX_)                                                                                                   \
    do {                                                                                                                   \
        PCS_CUDA(cudaFuncSetAttribute(k_linear_tc_wgrad<AG_, AX_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t.smem_bytes)); \
        k_linear_tc_wgrad<AG_, AX_><<<(unsigned)grid, kThreads, t.smem_bytes, s>>>(p);                                       \
    } while (0)
    if (ag && ax) PCS_WG(true, true);
    else if (ag) PCS_WG(true, false);
    else if (ax) PCS_WG(false, true);
    else PCS_WG(false, false);
#undef PCS_WG
    return check_launch("linear_tc_wgrad");
}

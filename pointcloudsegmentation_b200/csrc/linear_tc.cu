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
// accumulated in fp32 in TMEM: max-norm error 0.5-5e-6 of the result (stock fp32 GEMM: 0.2-1.2e-6).
//
// One persistent CTA per SM works on 128-row x BN-column tiles (BN <= 256), warp-specialised (20 warps):
//   warps 0-7   epilogue: warps e and e+4 drain TMEM lane quarter e%4 (rows 32*(e%4) .. +31), even / odd
//               16-column chunks, in one of three flavours chosen per launch (template parameter EPI):
//                 2  rows on a 16-byte pitch, no addend: tcgen05.ld (lane = row) -> bias / relu in registers -> ONE
//                    write into a 64-byte-swizzled staging tile -> TMA tensor store (cp.async.bulk.tensor.2d), the tensor
//                    map clips rows >= M and whole 16-byte units >= N; no transpose, no LDS, no STG;
//                 1  rows on a 16-byte pitch with an addend: shared-memory transpose, addend rows prefetched two chunks
//                    ahead in registers, 16-byte stores (the float4 that straddles column N element-wise);
//                 0  any alignment: shared-memory transpose, 4-byte loads / stores.
//               The bias lives in shared memory (staged once per CTA).  Two accumulator stages in TMEM (2 x 256
//               columns): the epilogue of tile i overlaps the MMAs of tile i+1.  These warps run with 120 registers
//               (setmaxnreg.inc) that warps 16-19 give up.
//   warps 8-15  A producers: read fp32 x rows straight from HBM (any K, any row stride: no TMA alignment
//               constraints), split into hi / lo in registers and store both into the 64-byte-swizzled
//               K-major layout tcgen05 reads (two 16-wide k-blocks per 128 bytes of a row, every lane busy in
//               both); loads run TWO 32-column units ahead in registers (HBM latency under load ~3000 clk).
//               One producer lane also issues ONE bulk async copy (cp.async.bulk, the TMA engine) per stage
//               for the B operand: the weight is pre-split and pre-swizzled into its shared-memory image by
//               k_linear_tc_prep_b (it is tiny and L2-resident), so a stage of B is one contiguous block.
//   warp 16     walks the pipeline as a whole warp; one elected lane (elect.sync) issues the 6 tcgen05.mma
//               (kind::tf32, M=128, N=BN, K=8; two k-steps x three split products) per 16-wide k-block and
//               commits them to the stage's "empty" mbarrier / the tile's "accumulator full" one.
//   warps 17-19 idle: registers are granted per warpgroup, so 17 warps cost the registers of 20 anyway; this
//               warpgroup drops to 40 registers per thread (setmaxnreg.dec) for the epilogue's benefit.
// Measured on B200 (tools/linear_tc_probe.py switches roles off through the act >> 8 bits): one 128x256x8 TF32 UMMA
// costs ~150 clk; a 16-wide k-block takes ~1360 clk for its 6 UMMAs (900 clk), so the tensor pipe is ~65 % busy on
// the wide layers and the rest is the producer -> mbarrier -> MMA -> commit round trip (~760 clk floor per k-block
// with everything else switched off).  The narrow layers (K <= 128) are bound by the EPILOGUE, and the epilogue by the
// L1 / shared-memory pipe and its own instruction stream, not by HBM (profiles/r3z_linear_tc_probe_stg.txt: storing into
// 1024 L2-resident rows costs the same as storing to HBM; ncu l1tex 65-79 %): hence flavour 2.
// [516607,128] -> 512: 429 -> 337 us, [516607,32] -> 96: 77 -> 56 us, -> 99 (odd, rows padded to a 16-byte pitch by the
// caller): 224 -> 81 us, [516607,131] -> 512 + addend + bias + relu: 1477 -> 656 us (profiles/r3t_*, r4b_*).
// Tried and dropped: 256-row tiles sharing one B
// stage between two UMMA halves (the single accumulator set serialises the epilogue), 128-byte-swizzled 32-wide
// k-blocks (2 stages only), issuing MMAs from a lane-0 branch (ptxas wraps each UTCHMMA in a waterfall loop),
// a lane -> 128-byte-chunk producer map (half the lanes idle in each 64-byte k-block), an epilogue that stages 32 x 32
// pieces as an image of the output modulo 16 bytes and stores aligned segments (two named barriers per piece: slower),
// issuing the next chunk's tcgen05.ld before the current chunk's stores (no gain: TMEM latency is not what the chunk waits for),
// taking the TMA flavour's chunks in pairs (one wait_group / proxy fence / commit per 32 columns: +-3 %, no gain).
#include <cuda.h>  // CUtensorMap (the encoder is fetched through cudaGetDriverEntryPoint: no libcuda link)

#include "common.cuh"
#include "tcgen05.cuh"

using namespace pcs;
using namespace pcs::tc;

namespace {
constexpr int BM = 128;             // rows per tile = UMMA M (cta_group::1): accumulator row i = TMEM lane i
constexpr int BK = 16;              // fp32 per k-block = 64 bytes = one SWIZZLE_64B row
constexpr int kEpiWarps = 8, kProdWarps = 8;
constexpr int kRowsPerWarp = BM / kProdWarps;  // 16
constexpr int kRegs4 = kRowsPerWarp / 4;       // float4 registers per lane and unit (both k-blocks)
// 8 epilogue + 8 producer warps + the MMA issuer, rounded up to whole warpgroups: registers are granted per 4 warps anyway
// (17 warps get the 96 registers per thread of 20), and the fifth warpgroup — the issuer and three idle warps — hands
// most of its registers to the epilogue warps (setmaxnreg), whose addend / accumulator sets do not fit into 96
constexpr int kThreads = (kEpiWarps + kProdWarps + 4) * 32;  // 640
constexpr int kRegsLaunch = 96, kRegsIssuer = 40, kRegsEpilogue = 120;
static_assert(4 * (kRegsLaunch - kRegsIssuer) >= kEpiWarps * (kRegsEpilogue - kRegsLaunch), "the epilogue's extra registers come from the issuer's warpgroup");
constexpr int kMaxStages = 6;
constexpr int kATile = BM * BK * 4;  // 8192 bytes (one of hi / lo)
constexpr int kStgPitch = 20;        // floats: 16 columns + 4 (conflict-free 16-byte row writes)
constexpr int kStgBytes = kEpiWarps * 32 * kStgPitch * 4;  // 20480
constexpr int kBarBytes = 256;
constexpr int kBiasCache = 1024;     // floats: the bias of all NT * BN columns, staged once per CTA (a per-chunk __ldg sat on
                                     // the epilogue's critical path and missed L1 behind the streaming stores)
constexpr int kSmemMax = 232448;  // 227 KB
constexpr uint32_t kTmemCols = 512;

// K-major, SWIZZLE_64B operand tile: rows of 64 bytes, 8-row atoms 512 bytes apart (SBO), LBO unused (=1),
// descriptor version 1 (sm_100), layout type 4 (cute::UMMA::SmemDescriptor, mma_sm100_desc.hpp)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | (32ull << 32) | (1ull << 46) | (4ull << 61);
}
// byte offset of (row r, 16-byte chunk c) in such a tile: chunk index XOR (r / 2) % 4  (Swizzle<2,4,3>)
__device__ __forceinline__ int sw64(int r, int c) { return r * 64 + ((c ^ ((r >> 1) & 3)) << 4); }

struct LinParams {
    const float *x;
    int64_t ldx;
    int M, K;
    const float *img;  // [NT][KB][2][BN][16] swizzled hi / lo image of B
    int N, BN, NT, KB;  // KB even: two k-blocks per 128 bytes of an x row
    const float *bias;
    const float *add;
    int64_t ldadd;
    float *y;
    int64_t ldy;
    int act;  // 0 none, 1 relu
    int vec_out;  // 16-byte row stores possible
    int stages, num_tiles;
    int stg_bytes;  // epilogue staging: kStgBytes, or 32 KB when the TMA flavour can double-buffer its 2 KB tiles
    int dbg;  // development switches (tools/linear_tc_probe.py; act >> 8): 1 no B copy, 4 no output, 8 no x loads, 16 no proxy fence,
              // 64 no epilogue chunks, 128 sums but no stores (flavour 1), 256 stores into 1024 rows (flavour 1); host side: 512 no
              // TMA flavour, 1024 one TMA staging tile per warp
};

// ---- weight image ----------------------------------------------------------------------------------
// img[nt][kb][h][n][16 floats]: exactly the bytes a shared-memory stage of B holds (64-byte rows, 16-byte
// chunks XOR-swizzled by (row / 2) % 4), zero padded in n and k, h = 0 hi / 1 lo.  B[n,k] = w[n*sn + k*sk].
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
        const int off = (sw64(n_in, k_in >> 2) >> 2) | (k_in & 3);
        float *base = img + ((int64_t)(nt * KB + kb) * 2) * BN * BK;
        base[off] = __uint_as_float(hi);
        base[BN * BK + off] = __uint_as_float(lo);
    }
}

// The float4 of an epilogue lane that straddles column N (rows 16-byte aligned, N % 4 != 0): its nc < 4 valid elements one
// by one.  o[i] = accumulator (the addend prefetch skips straddling pieces), rows i * 8 < nrow.  Cold path, kept out of the
// caller's register allocation.
// (The four float4 by value: a reference would put the caller's array into local memory.)
__device__ __noinline__ void store_straddling(float4 o0, float4 o1, float4 o2, float4 o3, int nc, int nrow, float *yp, int64_t ldy,
                                              const float *bp, const float *ap, int64_t ldadd, bool relu) {
    const float4 o[4] = {o0, o1, o2, o3};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (i * 8 >= nrow) continue;
        const float e[3] = {o[i].x, o[i].y, o[i].z};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            if (j >= nc) continue;
            float u = e[j];
            if (bp) u += bp[j];
            if (ap) u += __ldg(ap + (int64_t)(i * 8) * ldadd + j);
            if (relu) u = fmaxf(u, 0.f);
            yp[(int64_t)(i * 8) * ldy + j] = u;
        }
    }
}

// ---- main kernel -----------------------------------------------------------------------------------
// VEC: 16-byte loads of the x rows.  EPI: how the y rows leave — 0: 4-byte stores (any alignment), 1: 16-byte stores (rows
// on a 16-byte pitch; addend prefetched), 2: TMA tensor stores (16-byte pitch, no addend): the accumulator chunk is
// written ONCE into a 64-byte-swizzled staging tile, lane = row as it comes out of TMEM, and the TMA engine moves it —
// no transpose through shared memory, no LDS, no STG.  The narrow layers (K <= 128) are bound by the L1 / shared-memory
// pipe (ncu: l1tex 65-79 % busy; storing into 1024 L2-resident rows instead of HBM changes nothing, dropping the STG
// instructions does): per 512 bytes of output flavour 1 costs 4 (STS) + 4 (LDS) + 8 (STG, 64-byte row pieces) wavefronts,
// flavour 2 costs 4 + 1 (bias broadcast).
template <bool VEC, int EPI>
__global__ void __launch_bounds__(kThreads, 1)
k_linear_tc(const LinParams p, const __grid_constant__ CUtensorMap tmap_y) {
    constexpr bool EV = EPI == 1;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t *sm = smem_raw + (((raw + 1023u) & ~1023u) - raw);  // swizzle atoms need their natural alignment
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int stage_bytes = 2 * kATile + p.BN * 128;
    float *stg = reinterpret_cast<float *>(sm + p.stages * stage_bytes);
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<uint8_t *>(stg) + p.stg_bytes);
    const uint32_t bar0 = smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (kMaxStages + s); };
    auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + a); };
    auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kMaxStages + 2 + a); };
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * kMaxStages + 4);
    float *sbias = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(bars) + kBarBytes);
    const bool bias_cached = p.bias != nullptr && p.NT * p.BN <= kBiasCache;
    if (bias_cached)
        for (int i = threadIdx.x; i < p.NT * p.BN; i += kThreads) sbias[i] = i < p.N ? __ldg(p.bias + i) : 0.f;
    // TMEM: two accumulator stages of 256 columns each (BN <= 256): the epilogue of tile i overlaps the MMAs of i+1
    constexpr int acc_cols = 256;

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
        const int pw = warp - kEpiWarps;  // rows kRowsPerWarp*pw .. of the tile
        const int units = p.KB >> 1;      // 32 columns (two k-blocks) per unit
        // three register sets in rotation, indexed by compile-time constants (the loop below is unrolled by 3): a
        // cur = nxt; nxt = nx2 register shuffle makes every load complete within the iteration that issued it
        float4 R[3][kRegs4];
        // Register layout: r[2*i + h] = this lane's piece of row group i, k-block h of the unit, so that BOTH
        // store phases of a unit run with all 32 lanes active (a lane -> 128-byte-chunk map leaves half of the
        // lanes idle in each 64-byte k-block and doubles the producers' instruction count — they are the
        // critical warps, ncu).
        // lane -> (row lane/4 of 8, 16-byte chunk lane%4 of the 64 bytes of k-block h): one 16-byte load when the
        // row pieces are 16-byte aligned (VEC), four 4-byte loads otherwise (odd K such as 643, 131, 35) — the
        // shared-memory stores are 16 bytes wide either way; the swizzled offsets do not depend on tile or unit
        int off[kRowsPerWarp / 8];
#pragma unroll
        for (int i = 0; i < kRowsPerWarp / 8; ++i) off[i] = sw64(pw * kRowsPerWarp + i * 8 + (lane >> 2), lane & 3);
        auto load = [&](float4 (&r)[kRegs4], int tile, int unit) {
            const int row0 = (tile / p.NT) * BM + pw * kRowsPerWarp + (lane >> 2);
            const int col0 = unit * 32 + (lane & 3) * 4;
            const float *x0 = p.x + (int64_t)row0 * p.ldx + col0;
#pragma unroll
            for (int i = 0; i < kRowsPerWarp / 8; ++i) {
                const bool rok = row0 + i * 8 < p.M;
                const float *xr = x0 + (int64_t)(i * 8) * p.ldx;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int col = col0 + h * 16;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (rok && col < p.K) {
                        if (VEC) v = __ldg(reinterpret_cast<const float4 *>(xr + h * 16));
                        else if (col + 3 < p.K) v = make_float4(__ldg(xr + h * 16), __ldg(xr + h * 16 + 1), __ldg(xr + h * 16 + 2), __ldg(xr + h * 16 + 3));
                        else {
                            v.x = __ldg(xr + h * 16);
                            if (col + 1 < p.K) v.y = __ldg(xr + h * 16 + 1);
                            if (col + 2 < p.K) v.z = __ldg(xr + h * 16 + 2);
                        }
                    }
                    r[2 * i + h] = v;
                }
            }
        };
        // (tile, unit) sequence of this CTA; loads run TWO units (four k-blocks, ~6000 clk of MMA) ahead of the
        // stores in registers: one unit ahead the HBM latency under load (~3000 clk) was still exposed
        auto advance = [&](int &t, int &u) { if (++u == units) { u = 0; t += gridDim.x; } };
        int tile = blockIdx.x, unit = 0;
        uint32_t s = 0, ph = 0;
#pragma unroll
        for (int q = 0; q < 3; ++q)
#pragma unroll
            for (int i = 0; i < kRegs4; ++i) R[q][i] = make_float4(0.f, 0.f, 0.f, 0.f);
        int t1 = tile, u1 = 0;
        advance(t1, u1);
        int t2 = t1, u2 = u1;
        advance(t2, u2);
        if (!(p.dbg & 8)) {
            if (tile < p.num_tiles) load(R[0], tile, 0);
            if (t1 < p.num_tiles) load(R[1], t1, u1);
        }
        while (tile < p.num_tiles) {
#pragma unroll
          for (int q = 0; q < 3; ++q) {
            if (tile >= p.num_tiles) break;
            float4 (&cur)[kRegs4] = R[q];
            if (t2 < p.num_tiles && !(p.dbg & 8)) load(R[(q + 2) % 3], t2, u2);  // in flight while two units are split and stored
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int kb = unit * 2 + h;
                mbar_wait<20>(empty_bar((int)s), ph ^ 1u);
                uint8_t *a_hi = sm + s * stage_bytes, *a_lo = a_hi + kATile;
                if (pw == 0 && lane == 0) {
                    const uint32_t bytes = (uint32_t)p.BN * 128u;
                    const int nt = tile % p.NT;
                    const float *src = p.img + ((int64_t)(nt * p.KB + kb) * 2) * p.BN * BK;
                    if (p.dbg & 1) mbar_arrive(full_bar(s));
                    else {
                        mbar_arrive_expect_tx(full_bar(s), bytes);
                        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                     ::"r"(smem_u32(a_hi + 2 * kATile)), "l"(src), "r"(bytes), "r"(full_bar(s)) : "memory");
                    }
                }
#pragma unroll
                for (int i = 0; i < kRowsPerWarp / 8; ++i) {
                    const float4 v = cur[2 * i + h];
                    uint4 hh, ll;
                    hh.x = to_tf32(v.x); hh.y = to_tf32(v.y); hh.z = to_tf32(v.z); hh.w = to_tf32(v.w);
                    // lo is stored unrounded: the tensor core reads the top 19 bits (truncation, |err| < 2^-21 |x|)
                    ll.x = __float_as_uint(v.x - __uint_as_float(hh.x)); ll.y = __float_as_uint(v.y - __uint_as_float(hh.y));
                    ll.z = __float_as_uint(v.z - __uint_as_float(hh.z)); ll.w = __float_as_uint(v.w - __uint_as_float(hh.w));
                    *reinterpret_cast<uint4 *>(a_hi + off[i]) = hh;
                    *reinterpret_cast<uint4 *>(a_lo + off[i]) = ll;
                }
                if (!(p.dbg & 16)) fence_async_smem();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
                __syncwarp();
                if (lane == 0) mbar_arrive(full_bar((int)s));
                if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
            }
            tile = t1; unit = u1; t1 = t2; u1 = u2;
            advance(t2, u2);
          }
        }
    } else if (warp >= kEpiWarps + kProdWarps) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsIssuer) : "memory");
        if (warp == kEpiWarps + kProdWarps)
        // ================= MMA issuer: the whole warp walks the loop, one elected lane issues =================
        // tcgen05.mma issue is in order and the queue is shallow: whatever this warp does between the last MMA of one
        // k-block and the first of the next is a bubble in the tensor pipe (measured: ~500 clk per k-block with
        // divisions and descriptor builds in the loop).  So the loop carries only counters and four 64-bit adds.
        {
            // cute::UMMA::InstrDescriptor: c_format F32 (1<<4), a/b_format TF32 (2<<7, 2<<10), K-major both, N>>3, M>>4
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((128u >> 4) << 24);
            const uint32_t base0 = smem_u32(sm);
            const uint64_t dah0 = umma_desc(base0), dal0 = umma_desc(base0 + kATile);
            const uint64_t dbh0 = umma_desc(base0 + 2 * kATile), dbl0 = umma_desc(base0 + 2 * kATile + p.BN * 64);
            const uint64_t dstep = (uint64_t)(stage_bytes >> 4);  // next stage in the descriptors' (addr >> 4) field
            uint32_t s = 0, ph = 0, tl = 0;
            uint64_t soff = 0;
            for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x, ++tl) {
                const uint32_t a = tl & 1u;
                mbar_wait<100>(tempty_bar((int)a), ((tl >> 1) & 1u) ^ 1u);  // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t d = tmem_base + a * (uint32_t)acc_cols;
                for (int kb = 0; kb < p.KB; ++kb) {
                    mbar_wait<0>(full_bar((int)s), ph);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint64_t dah = dah0 + soff, dal = dal0 + soff, dbh = dbh0 + soff, dbl = dbl0 + soff;
                        // UMMA K = 8 tf32 = 32 bytes: +2 in the (addr >> 4) field per k-step
                        tc_mma_tf32(d, dal, dbh, idesc, kb > 0 ? 1u : 0u);
                        tc_mma_tf32(d, dah, dbl, idesc, 1u);
                        tc_mma_tf32(d, dah, dbh, idesc, 1u);
                        tc_mma_tf32(d, dal + 2, dbh + 2, idesc, 1u);
                        tc_mma_tf32(d, dah + 2, dbl + 2, idesc, 1u);
                        tc_mma_tf32(d, dah + 2, dbh + 2, idesc, 1u);
                        tc_commit(empty_bar((int)s));  // stage reusable once these MMAs have read it
                        if (kb == p.KB - 1) tc_commit(tfull_bar((int)a));  // accumulators complete
                    }
                    __syncwarp();
                    if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; soff = 0; } else soff += dstep;
                }
            }
        }
    } else {
        // ================= epilogue: TMEM -> registers -> smem transpose -> coalesced rows =================
        // The 8 epilogue warps never wait on the narrow layers (K <= 128): they ARE the critical path there, and what
        // bounds them is their own instruction stream (ncu source view, [516607,32] -> 96: 700 instructions per tile and
        // warp, no spin).  So everything that depends only on the tile is computed once per tile, the parameters live in
        // registers, and the store flavour EV is a template parameter.
        asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsEpilogue) : "memory");  // waits for the issuer group's dec
        const int q = warp & 3, par = warp >> 2;  // TMEM lane quarter; even / odd 16-column chunks
        const int M = p.M, N = p.N, BN = p.BN, NT = p.NT;  // (constant-bank operands where the compiler sees fit)
        const int64_t ldy = p.ldy, ldadd = p.ldadd;
        const float *const bias = p.bias;
        const float *const add = p.add;
        float *const y = p.y;
        const bool relu = p.act == 1, no_store = (p.dbg & 4) != 0;
        const int nch = (p.dbg & 64) ? 0 : BN / 16;
        const int MT = (M + BM - 1) / BM;
        const int smt = (int)gridDim.x / NT, snt = (int)gridDim.x % NT;  // tile += gridDim.x without a division per tile
        auto next_tile = [&](int &mt, int &nt) { mt += smt; nt += snt; if (nt >= NT) { nt -= NT; ++mt; } };
        // lane -> (row lr + 8 i, columns lc .. lc + 3), i < 4, of a 32-row x 16-column chunk; 4-byte flavour: (row lr + 2 i,
        // column lc), i < 16
        const int lr = EV ? lane >> 2 : lane >> 4;
        const int lc = EV ? (lane & 3) * 4 : (lane & 15);
        float *const stg_wr = stg + warp * 32 * kStgPitch + lane * kStgPitch;
        const float *const stg_rd = stg + warp * 32 * kStgPitch + lr * kStgPitch + lc;
        const uint32_t tq = tmem_base + ((uint32_t)(q * 32) << 16);
        int mt = (int)blockIdx.x / NT, nt = (int)blockIdx.x % NT;
        uint32_t tl = 0;
        if constexpr (EPI == 2) {
            // staging tiles of 32 rows x 64 bytes, SWIZZLE_64B (512-byte aligned); two per warp where shared memory allows, so
            // that a chunk is written while the TMA engine still reads the previous one
            const bool two = p.stg_bytes >= kEpiWarps * 4096;
            uint8_t *const buf0 = reinterpret_cast<uint8_t *>(stg) + warp * (two ? 4096 : 2048);
            uint32_t ib = 0;
            const uint64_t tm = reinterpret_cast<uint64_t>(&tmap_y);
            const float *const bsrc = bias_cached ? sbias : bias;
            const int N4 = N & ~3;  // the tensor map ends here: TMA clips in 16-byte units and would write up to column N4 + 4
            for (; mt < MT; next_tile(mt, nt), ++tl) {
                const uint32_t a = tl & 1u;
                mbar_wait<200>(tfull_bar((int)a), (tl >> 1) & 1u);
                tc_fence_after();
                const int row0 = mt * BM + q * 32;
                for (int c = par; c < nch; c += 2) {
                    const int col = nt * BN + c * 16;
                    uint32_t v[16];
                    tmem_ld16(tq + a * (uint32_t)acc_cols + (uint32_t)(c * 16), v);
                    if (no_store || col >= N) continue;  // (warp-uniform)
                    float f[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) f[j] = __uint_as_float(v[j]);
                    if (bias) {  // the same 16 columns for every lane (= row): broadcast reads
                        if (bias_cached) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const float4 b = *reinterpret_cast<const float4 *>(bsrc + col + 4 * j);
                                f[4 * j] += b.x; f[4 * j + 1] += b.y; f[4 * j + 2] += b.z; f[4 * j + 3] += b.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                if (col + j < N) f[j] += __ldg(bsrc + col + j);
                        }
                    }
                    if (relu) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) f[j] = fmaxf(f[j], 0.f);
                    }
                    uint8_t *const buf = buf0 + ib * 2048;
                    if (lane == 0) {  // the store that last used this tile has read it
                        if (two) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                        else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                    }
                    __syncwarp();
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        *reinterpret_cast<float4 *>(buf + sw64(lane, j)) = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                    fence_async_smem();  // generic-proxy stores -> visible to the TMA engine's async-proxy reads
                    __syncwarp();
                    if (lane == 0 && row0 < M && col < N4) {  // rows >= M and columns >= N4 of the box are clipped by the tensor map
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(tm), "r"(smem_u32(buf)), "r"(col), "r"(row0) : "memory");
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                    if (two) ib ^= 1u;
                    if (N4 < N && col + 16 > N4 && row0 + lane < M) {  // the <= 3 columns behind the last whole 16 bytes of a row
                        float *yr = y + (int64_t)(row0 + lane) * ldy + col;
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            if (col + j >= N4 && col + j < N) yr[j] = f[j];
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty_bar((int)a));
            }
            if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // shared memory outlives the stores
        } else if (par >= nch) {
            // no chunk for this warp (BN = 16, or the debug switch): keep the accumulator hand-shake going
            for (; mt < MT; next_tile(mt, nt), ++tl) {
                const uint32_t a = tl & 1u;
                mbar_wait<200>(tfull_bar((int)a), (tl >> 1) & 1u);
                tc_fence_after();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(tempty_bar((int)a));
            }
        } else {
            // The addend rows do not depend on the accumulators: they are loaded TWO chunks ahead (across tile
            // boundaries, so also while this warp waits for the MMAs of its next tile).  Loaded inside the chunk, one
            // chunk's 4 x 16 B per lane was all an SM had in flight (16 KB): the 131 -> 512 classifier layer with its
            // 1 GB addend ran 1.5 ms, four times its HBM time.  Three register sets in rotation with compile-time
            // indices, as in the producers.  (EV only: the 4-byte flavour loads its addend where it uses it.)
            const bool pre = EV && add != nullptr;
            float4 AD[3][4];
            int pmt = mt, pnt = nt, pc = par;  // the next chunk to prefetch
            auto prefetch = [&](float4 (&r)[4]) {
                if (pmt < MT) {
                    const int row0 = pmt * BM + q * 32 + lr, col = pnt * BN + pc * 16 + lc;
                    const float *ap = add + (int64_t)row0 * ldadd + col;
                    const int nrow = M - row0;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        r[i] = (col + 3 < N && i * 8 < nrow) ? __ldg(reinterpret_cast<const float4 *>(ap + (int64_t)(i * 8) * ldadd))
                                                             : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                pc += 2;
                if (pc >= nch) { pc = par; next_tile(pmt, pnt); }
            };
#pragma unroll
            for (int k = 0; k < 3; ++k)
#pragma unroll
                for (int i = 0; i < 4; ++i) AD[k][i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (pre) { prefetch(AD[0]); prefetch(AD[1]); }
            int c = par;
            // per tile: rows / columns left from this lane's first element, its output (addend, bias) pointers
            int nrow = 0, ncol = 0;
            float *yp = y;
            const float *ap = add, *bp = bias;
            uint32_t tacc = tq;
            while (mt < MT) {
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    if (mt >= MT) break;
                    if (pre) prefetch(AD[(k + 2) % 3]);
                    if (c == par) {  // first chunk of a tile
                        const int row0 = mt * BM + q * 32 + lr, col0 = nt * BN + lc;
                        nrow = M - row0;
                        ncol = N - col0;
                        yp = y + (int64_t)row0 * ldy + col0;
                        if (!EV && add) ap = add + (int64_t)row0 * ldadd + col0;
                        if (bias) bp = (bias_cached ? sbias : bias) + col0;
                        const uint32_t a = tl & 1u;
                        tacc = tq + a * (uint32_t)acc_cols;
                        mbar_wait<200>(tfull_bar((int)a), (tl >> 1) & 1u);
                        tc_fence_after();
                    }
                    const int cc = c * 16;
                    {
                        uint32_t v[16];
                        tmem_ld16(tacc + (uint32_t)cc, v);
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<uint4 *>(stg_wr + j * 4) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    }
                    __syncwarp();
                    if (EV) {
                        float4 (&o)[4] = AD[k];  // the sums are formed in the addend's registers
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 v = *reinterpret_cast<const float4 *>(stg_rd + i * 8 * kStgPitch);
                            if (pre) { o[i].x += v.x; o[i].y += v.y; o[i].z += v.z; o[i].w += v.w; }
                            else o[i] = v;
                        }
                        if (no_store) {
                        } else if (cc + 3 < ncol) {  // the lane's four columns are inside the row
                            if (bias) {
                                const float4 b = *reinterpret_cast<const float4 *>(bp + cc);
#pragma unroll
                                for (int i = 0; i < 4; ++i) { o[i].x += b.x; o[i].y += b.y; o[i].z += b.z; o[i].w += b.w; }
                            }
                            if (relu) {
#pragma unroll
                                for (int i = 0; i < 4; ++i) { o[i].x = fmaxf(o[i].x, 0.f); o[i].y = fmaxf(o[i].y, 0.f); o[i].z = fmaxf(o[i].z, 0.f); o[i].w = fmaxf(o[i].w, 0.f); }
                            }
                            float *yc = yp + cc;
                            if (p.dbg & 256) yc = y + (int64_t)((mt * BM + q * 32 + lr) & 1023) * ldy + nt * BN + lc + cc;  // L2-resident target
                            if (p.dbg & 128) {  // the sums, but no store
                                if (o[0].x + o[1].y + o[2].z + o[3].w == 123.456f) yc[0] = 0.f;
                            } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i)
                                if (i * 8 < nrow) *reinterpret_cast<float4 *>(yc + (int64_t)(i * 8) * ldy) = o[i];
                            }
                        } else if (cc < ncol) {  // they straddle column N (N % 4 != 0 on 16-byte aligned rows): out of line
                            const int row0 = mt * BM + q * 32 + lr, col = nt * BN + lc + cc;
                            store_straddling(o[0], o[1], o[2], o[3], ncol - cc, nrow, yp + cc, ldy, bias ? bp + cc : nullptr,
                                             add ? add + (int64_t)row0 * ldadd + col : nullptr, ldadd, relu);
                        }
                    } else {
                        float o[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) o[i] = stg_rd[i * 2 * kStgPitch];
                        if (!no_store && cc < ncol) {
                            const float b = bias ? bp[cc] : 0.f;
                            float *yc = yp + cc;
                            const float *ac = ap + cc;
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                if (i * 2 < nrow) {
                                    float u = o[i] + b;
                                    if (add) u += __ldg(ac + (int64_t)(i * 2) * ldadd);
                                    if (relu) u = fmaxf(u, 0.f);
                                    yc[(int64_t)(i * 2) * ldy] = u;
                                }
                            }
                        }
                    }
                    __syncwarp();
                    c += 2;
                    if (c >= nch) {  // that was the tile's last chunk for this warp: hand the accumulator back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(tempty_bar((int)(tl & 1u)));
                        ++tl;
                        c = par;
                        next_tile(mt, nt);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// Tensor map of y[M, N] (N and the row pitch ldy multiples of 4 floats; base 16-byte aligned) for the epilogue's TMA stores:
// boxes of 32 rows x 16 columns, 64-byte swizzle.  (The engine clips a box at the tensor's edge in 16-byte units: with
// N = 5 it wrote columns 5..7 too, so the caller passes N rounded DOWN and stores the last <= 3 columns itself.)  cuTensorMapEncodeTiled is a host-only driver call (no context work): fetched
// once through the runtime, so the library does not link libcuda.  false: not available / not encodable (the caller keeps
// the 16-byte store flavour).
bool make_tmap_y(CUtensorMap *tm, float *y, int64_t M, int64_t N, int64_t ldy) {
    using Encode = CUresult (*)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static Encode encode = [] {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess)
            fn = nullptr;
        (void)cudaGetLastError();
        return reinterpret_cast<Encode>(fn);
    }();
    if (!encode || ldy * 4 >= (1ll << 40)) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M};
    const cuuint64_t strides[1] = {(cuuint64_t)ldy * 4};
    const cuuint32_t box[2] = {16, 32}, estr[2] = {1, 1};
    return encode(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, y, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct Tiling {
    int NT, BN, KB, stages;
    size_t img_bytes, smem_bytes;
};
Tiling make_tiling_tc(int64_t N, int64_t K) {
    Tiling t;
    t.NT = (int)((N + 255) / 256);
    const int64_t per = (N + t.NT - 1) / t.NT;
    t.BN = (int)((per + 15) / 16 * 16);
    t.KB = (int)((K + 31) / 32) * 2;
    const int stage_bytes = 2 * kATile + t.BN * 128;
    const int avail = kSmemMax - kStgBytes - kBarBytes - kBiasCache * 4 - 1024;
    t.stages = avail / stage_bytes;
    if (t.stages > kMaxStages) t.stages = kMaxStages;
    t.smem_bytes = (size_t)t.stages * stage_bytes + kStgBytes + kBarBytes + kBiasCache * 4 + 1024;
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
    p.vec_out = (ldy % 4 == 0) && aligned(y, 16) && (!bias || aligned(bias, 16)) &&
                (!addend || (ldadd % 4 == 0 && aligned(addend, 16)));
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
    // 16-byte stores need aligned rows, not N % 4 == 0 (the float4 that straddles column N is written element-wise);
    // without an addend the rows leave through the TMA engine
    int epi = p.vec_out ? 1 : 0;
    CUtensorMap tmap;
    memset(&tmap, 0, sizeof(tmap));
    if (epi == 1 && !addend && !(dbg & 512) && N >= 4 && make_tmap_y(&tmap, y, M, N & ~3ll, ldy)) epi = 2;
    size_t smem_bytes = t.smem_bytes;
    p.stg_bytes = kStgBytes;
    if (epi == 2 && !(dbg & 1024) && smem_bytes + (kEpiWarps * 4096 - kStgBytes) <= (size_t)kSmemMax) {  // room for two staging tiles per warp
        p.stg_bytes = kEpiWarps * 4096;
        smem_bytes += kEpiWarps * 4096 - kStgBytes;
    }
    auto launch = [&](auto kern) -> int {
        // the register hand-over inside the kernel (setmaxnreg) is sized for kRegsLaunch registers per thread at launch: with
        // fewer the epilogue warps would wait for registers nobody releases — refuse loudly instead
        static int regs = -1;  // (one per kernel instantiation: the lambda is generic)
        if (regs < 0) {
            cudaFuncAttributes fa;
            PCS_CUDA(cudaFuncGetAttributes(&fa, kern));
            regs = fa.numRegs;
        }
        PCS_REQUIRE(regs == kRegsLaunch, "linear_tc: kernel built with %d registers per thread, the register hand-over needs %d", regs,
                    kRegsLaunch);
        PCS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
        kern<<<grid, kThreads, smem_bytes, s>>>(p, tmap);
        return 0;
    };
    int e;
    if (vec) e = epi == 2 ? launch(k_linear_tc<true, 2>) : epi == 1 ? launch(k_linear_tc<true, 1>) : launch(k_linear_tc<true, 0>);
    else e = epi == 2 ? launch(k_linear_tc<false, 2>) : epi == 1 ? launch(k_linear_tc<false, 1>) : launch(k_linear_tc<false, 0>);
    if (e) return e;
    return check_launch("linear_tc");
}

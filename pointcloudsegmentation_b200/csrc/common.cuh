// common.cuh — shared device/host helpers for libpcseg_ops.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include <initializer_list>
#include "../../include/pcseg_ops.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libpcseg_ops targets sm_100a (B200) only"
#endif

namespace pcs {

// ---- error / bookkeeping ------------------------------------------------------------------
extern thread_local char g_err[512];
extern std::atomic<int64_t> g_launches;

// tuning knobs, set through pcs_set_option (common.cu)
namespace opt {
extern std::atomic<int64_t> scatter_walk_min_c_diff, scatter_walk_min_c_diff_odd, scatter_walk_min_c_copy,
    scatter_walk_min_c_copy_odd, mlp_tensor_cores, pool_wgrad_staged, walk_rows_narrow, seg_stream, flat_small_c, elim_tile;
}

int fail(const char *fmt, ...);
int check_launch(const char *what);

#define PCS_REQUIRE(cond, ...)                  \
    do {                                        \
        if (!(cond)) return pcs::fail(__VA_ARGS__); \
    } while (0)

#define PCS_CUDA(expr)                                                                   \
    do {                                                                                 \
        cudaError_t _e = (expr);                                                         \
        if (_e != cudaSuccess) return pcs::fail("%s: %s", #expr, cudaGetErrorString(_e)); \
    } while (0)

static inline cudaStream_t as_stream(pcs_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

static inline int next_pow2(int64_t v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}
static inline int ilog2(int p) {
    int s = 0;
    while ((1 << s) < p) ++s;
    return s;
}
static inline bool aligned(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Row tiling shared by every feature-moving kernel: a row of C floats is covered by CH chunks of
// G lanes x V floats (G a power of two <= 32), channel index on the fast lane axis so that a
// row access is one or more full 128-byte lines (the opposite of the reference's heuristic,
// TFNeighborKernel.cu:266-280, which puts the POINT index on threadIdx.x).
struct RowTiling {
    int V;       // floats per lane: 4, 2 or 1
    int gshift;  // log2(G); -1 = exact tiling (thread t -> item t / CH, column (t % CH) * V)
    int CH;      // chunks per row (exact tiling: lanes per item = C / V)
};
static inline RowTiling make_tiling(int64_t C, bool can4, bool can2) {
    RowTiling t;
    t.V = (can4 && C % 4 == 0) ? 4 : ((can2 && C % 2 == 0) ? 2 : 1);
    int64_t lanes = (C + t.V - 1) / t.V;
    int G = lanes >= 32 ? 32 : next_pow2(lanes);
    t.gshift = ilog2(G);
    t.CH = (int)((lanes + G - 1) / G);
    // Widths whose lane count is not a power of two (C = 3, 12, 35, 48, 96, 131 ...): power-of-two groups
    // would idle up to half of the lanes (C = 35 -> 64 slots, C = 48 -> 16 slots for 12 lanes); any
    // thread -> (item, column) map is coalesced, so use the exact one and keep every lane busy.
    // (Not for < 8 vector lanes per row, e.g. C = 12: 3-lane groups straddle warps and the row writers
    // lose more than the idle 4th lane costs — measured, profiles/README.md.)
    if ((int64_t)G * t.CH != lanes && (t.V == 1 || lanes >= 8)) { t.gshift = -1; t.CH = (int)lanes; }
    return t;
}

// ---- device vector helpers ------------------------------------------------------------------
template <int V> struct Vec { float v[V]; };

template <int V> __device__ __forceinline__ Vec<V> ld(const float *p);
template <> __device__ __forceinline__ Vec<1> ld<1>(const float *p) { Vec<1> r; r.v[0] = __ldg(p); return r; }
template <> __device__ __forceinline__ Vec<2> ld<2>(const float *p) {
    float2 t = __ldg(reinterpret_cast<const float2 *>(p));
    Vec<2> r; r.v[0] = t.x; r.v[1] = t.y; return r;
}
template <> __device__ __forceinline__ Vec<4> ld<4>(const float *p) {
    float4 t = __ldg(reinterpret_cast<const float4 *>(p));
    Vec<4> r; r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w; return r;
}
template <int V> __device__ __forceinline__ void st(float *p, const Vec<V> &a);
template <> __device__ __forceinline__ void st<1>(float *p, const Vec<1> &a) { *p = a.v[0]; }
template <> __device__ __forceinline__ void st<2>(float *p, const Vec<2> &a) {
    *reinterpret_cast<float2 *>(p) = make_float2(a.v[0], a.v[1]);
}
template <> __device__ __forceinline__ void st<4>(float *p, const Vec<4> &a) {
    *reinterpret_cast<float4 *>(p) = make_float4(a.v[0], a.v[1], a.v[2], a.v[3]);
}
template <int V> __device__ __forceinline__ Vec<V> vzero() {
    Vec<V> r;
#pragma unroll
    for (int k = 0; k < V; ++k) r.v[k] = 0.f;
    return r;
}
template <int V> __device__ __forceinline__ void vadd(Vec<V> &a, const Vec<V> &b) {
#pragma unroll
    for (int k = 0; k < V; ++k) a.v[k] += b.v[k];
}
template <int V> __device__ __forceinline__ Vec<V> vsub(const Vec<V> &a, const Vec<V> &b) {
    Vec<V> r;
#pragma unroll
    for (int k = 0; k < V; ++k) r.v[k] = a.v[k] - b.v[k];
    return r;
}

// Decompose the flat thread id into (row item, chunk, lane column).  Returns false when idle.
__device__ __forceinline__ bool row_item(int64_t n_items, int gshift, int CH, int V, int64_t C,
                                         int64_t &item, int &col) {
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gshift < 0) {  // exact tiling: CH lanes per item
        if (t < ((int64_t)1 << 31)) {
            const unsigned tt = (unsigned)t, cc = (unsigned)CH;
            const unsigned q = tt / cc;
            item = q;
            col = (int)(tt - q * cc) * V;
        } else {
            item = t / CH;
            col = (int)(t - item * CH) * V;
        }
        return item < n_items;
    }
    int64_t gid = t >> gshift;
    int lane = (int)(t & ((1 << gshift) - 1));
    item = gid / CH;
    int ch = (int)(gid - item * CH);
    col = ((ch << gshift) + lane) * V;
    return item < n_items && col < C;
}

static inline int64_t grid_for(int64_t n_items, const RowTiling &t, int block) {
    int64_t threads = t.gshift < 0 ? n_items * t.CH : ((n_items * t.CH) << t.gshift);
    return (threads + block - 1) / block;
}


// ---- storage types: fp32 or bf16 rows, fp32 arithmetic ---------------------------------------------
// Every feature mover is templated on the STORAGE type T; a lane owns V consecutive elements (V * sizeof(T)
// <= 16 bytes, one LDG/STG), converts to fp32 on load and rounds (RN-even) on store.  bf16 -> fp32 is a
// 16-bit shift, so copies and max/argmax are exact; sums accumulate in fp32 (north_star: "bf16x8 lanes").
typedef __nv_bfloat16 bf16;

template <typename T, int V> struct IO;
template <int V> struct IO<float, V> {
    static __device__ __forceinline__ Vec<V> ldg(const float *p) { return ld<V>(p); }
    static __device__ __forceinline__ Vec<V> lds(const float *p) {  // shared / generic pointer, no __ldg
        Vec<V> r;
        if constexpr (V == 4) { float4 t = *reinterpret_cast<const float4 *>(p); r.v[0] = t.x; r.v[1] = t.y; r.v[2] = t.z; r.v[3] = t.w; }
        else if constexpr (V == 2) { float2 t = *reinterpret_cast<const float2 *>(p); r.v[0] = t.x; r.v[1] = t.y; }
        else r.v[0] = *p;
        return r;
    }
    static __device__ __forceinline__ void stg(float *p, const Vec<V> &a) { st<V>(p, a); }
};
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t bf_pack(float lo, float hi) {
    __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);  // .x = lo (low half), .y = hi
    return *reinterpret_cast<uint32_t *>(&t);
}
template <int V> struct IO<bf16, V> {
    template <bool LDG> static __device__ __forceinline__ Vec<V> load(const bf16 *p) {
        Vec<V> r;
        if constexpr (V == 1) {
            const unsigned short u = LDG ? __ldg(reinterpret_cast<const unsigned short *>(p)) : *reinterpret_cast<const unsigned short *>(p);
            r.v[0] = __uint_as_float((uint32_t)u << 16);
        } else if constexpr (V == 2) {
            const uint32_t w = LDG ? __ldg(reinterpret_cast<const uint32_t *>(p)) : *reinterpret_cast<const uint32_t *>(p);
            r.v[0] = bf_lo(w); r.v[1] = bf_hi(w);
        } else if constexpr (V == 4) {
            const uint2 w = LDG ? __ldg(reinterpret_cast<const uint2 *>(p)) : *reinterpret_cast<const uint2 *>(p);
            r.v[0] = bf_lo(w.x); r.v[1] = bf_hi(w.x); r.v[2] = bf_lo(w.y); r.v[3] = bf_hi(w.y);
        } else {
            const uint4 w = LDG ? __ldg(reinterpret_cast<const uint4 *>(p)) : *reinterpret_cast<const uint4 *>(p);
            r.v[0] = bf_lo(w.x); r.v[1] = bf_hi(w.x); r.v[2] = bf_lo(w.y); r.v[3] = bf_hi(w.y);
            r.v[4] = bf_lo(w.z); r.v[5] = bf_hi(w.z); r.v[6] = bf_lo(w.w); r.v[7] = bf_hi(w.w);
        }
        return r;
    }
    static __device__ __forceinline__ Vec<V> ldg(const bf16 *p) { return load<true>(p); }
    static __device__ __forceinline__ Vec<V> lds(const bf16 *p) { return load<false>(p); }
    static __device__ __forceinline__ void stg(bf16 *p, const Vec<V> &a) {
        if constexpr (V == 1) *p = __float2bfloat16_rn(a.v[0]);
        else if constexpr (V == 2) *reinterpret_cast<uint32_t *>(p) = bf_pack(a.v[0], a.v[1]);
        else if constexpr (V == 4) *reinterpret_cast<uint2 *>(p) = make_uint2(bf_pack(a.v[0], a.v[1]), bf_pack(a.v[2], a.v[3]));
        else *reinterpret_cast<uint4 *>(p) = make_uint4(bf_pack(a.v[0], a.v[1]), bf_pack(a.v[2], a.v[3]),
                                                        bf_pack(a.v[4], a.v[5]), bf_pack(a.v[6], a.v[7]));
    }
};

// Largest power-of-two alignment (<= 16) shared by all the given pointers.
static inline int common_align(std::initializer_list<const void *> ps) {
    uintptr_t m = 0;
    for (const void *p : ps) m |= reinterpret_cast<uintptr_t>(p);
    int a = 16;
    while (a > 1 && (m & (uintptr_t)(a - 1))) a >>= 1;
    return a;
}

// make_tiling for a storage type of `es` bytes: V = the widest lane (<= 16 bytes, <= align) that divides C.
static inline RowTiling make_tiling_es(int64_t C, int es, int align) {
    RowTiling t;
    t.V = 1;
    for (int v = 16 / es; v >= 1; v >>= 1)
        if (C % v == 0 && v * es <= align) { t.V = v; break; }
    int64_t lanes = C / t.V;
    int G = lanes >= 32 ? 32 : next_pow2(lanes);
    t.gshift = ilog2(G);
    t.CH = (int)((lanes + G - 1) / G);
    if ((int64_t)G * t.CH != lanes && (t.V == 1 || lanes >= 8)) { t.gshift = -1; t.CH = (int)lanes; }
    return t;
}

// Dispatch on (storage type, lane width): defines `T` and `constexpr int V` for the statement(s) given.
#define PCS_TV_CASE(V_, ...) case V_: { constexpr int V = V_; __VA_ARGS__; } break;
#define PCS_DISPATCH_TV(IS_BF16_, VV_, ...)                                                         \
    do {                                                                                            \
        if (IS_BF16_) {                                                                             \
            using T = pcs::bf16;                                                                    \
            switch (VV_) { PCS_TV_CASE(8, __VA_ARGS__) PCS_TV_CASE(4, __VA_ARGS__) PCS_TV_CASE(2, __VA_ARGS__) \
                           default: { constexpr int V = 1; __VA_ARGS__; } break; }                  \
        } else {                                                                                    \
            using T = float;                                                                        \
            switch (VV_) { PCS_TV_CASE(4, __VA_ARGS__) PCS_TV_CASE(2, __VA_ARGS__)                    \
                           default: { constexpr int V = 1; __VA_ARGS__; } break; }                  \
        }                                                                                           \
    } while (0)

}  // namespace pcs

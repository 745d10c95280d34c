// common.cuh — shared device/host helpers for libpcseg_ops.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
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
    scatter_walk_min_c_copy_odd, mlp_tensor_cores;
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

}  // namespace pcs

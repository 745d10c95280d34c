// search_boxes.cuh — the two-level bounding boxes the culled searches share (search.cu: a18-a20 self searches;
// cross_search.cu: query-set-vs-source-set radius search and k-NN).  Included by exactly those translation units;
// everything here has internal linkage.
#pragma once
#include "common.cuh"
#include <math.h>

namespace {
using namespace pcs;

// Bounding box of every 32-point chunk [32g, 32g+32) of the whole array (SoA: 6 x n_chunks floats).
__global__ void __launch_bounds__(256)
k_chunk_boxes(const float *__restrict__ xyzs, int64_t pn, int64_t n_chunks, float *__restrict__ boxes) {
    const int64_t g = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (g >= n_chunks) return;
    const int64_t j = g * 32 + lane;
    const float inf = __int_as_float(0x7f800000);
    float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
    if (j < pn) {
#pragma unroll
        for (int k = 0; k < 3; ++k) lo[k] = hi[k] = __ldg(xyzs + j * 3 + k);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int off = 16; off > 0; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
        }
    if (lane < 3) boxes[lane * n_chunks + g] = lo[lane];
    else if (lane < 6) boxes[lane * n_chunks + g] = hi[lane - 3];
}

// Level-2 boxes: super-chunk s = union of the chunk boxes [32s, 32s+32).  Same SoA layout, n_super entries.
__global__ void __launch_bounds__(256)
k_super_boxes(const float *__restrict__ boxes, int64_t n_chunks, int64_t n_super, float *__restrict__ sboxes) {
    const int64_t sidx = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (sidx >= n_super) return;
    const int64_t g = sidx * 32 + lane;
    const float inf = __int_as_float(0x7f800000);
    float v[6] = {inf, inf, inf, -inf, -inf, -inf};
    if (g < n_chunks) {
#pragma unroll
        for (int k = 0; k < 6; ++k) v[k] = __ldg(boxes + k * n_chunks + g);
    }
#pragma unroll
    for (int k = 0; k < 6; ++k)
        for (int off = 16; off > 0; off >>= 1) {
            const float o = __shfl_xor_sync(0xffffffffu, v[k], off);
            v[k] = k < 3 ? fminf(v[k], o) : fmaxf(v[k], o);
        }
    if (lane < 6) {
        float out = v[0];
#pragma unroll
        for (int k = 1; k < 6; ++k) out = lane == k ? v[k] : out;
        sboxes[lane * n_super + sidx] = out;
    }
}
}  // namespace

static inline int64_t n_chunks_of(int64_t pn) { return (pn + 31) / 32; }
static inline int64_t n_super_of(int64_t pn) { return (n_chunks_of(pn) + 31) / 32; }
static inline size_t boxes_bytes(int64_t pn) {
    return align_up((size_t)(n_chunks_of(pn) + n_super_of(pn)) * 6 * sizeof(float), 256);
}

// sqrt(sq_max) * (1 + 1e-3): a chunk farther than this on one axis cannot hold a hit (fp32 rounding of
// the predicate is ~1e-7 relative).  NaN / negative sq_max => NaN => nothing is culled.
static float cull_radius(float sq_max) { return sqrtf(sq_max) * 1.001f; }

static int make_boxes(const float *xyzs, int64_t pn, void *workspace, size_t workspace_bytes, cudaStream_t s,
                      float **boxes, float **sboxes) {
    PCS_REQUIRE(workspace && workspace_bytes >= boxes_bytes(pn), "search: workspace too small");
    *boxes = static_cast<float *>(workspace);
    const int64_t nc = n_chunks_of(pn), ns = n_super_of(pn);
    *sboxes = *boxes + nc * 6;
    k_chunk_boxes<<<(unsigned)((nc * 32 + 255) / 256), 256, 0, s>>>(xyzs, pn, nc, *boxes);
    if (check_launch("chunk_boxes")) return 1;
    k_super_boxes<<<(unsigned)((ns * 32 + 255) / 256), 256, 0, s>>>(*boxes, nc, ns, *sboxes);
    return check_launch("super_boxes");
}


// neighbor_misc.cu — a10 LocationWeightFeatSumBackward (fp32 / bf16), a11 EliminateCenter, a12 ComputeVoxelIndex,
// a16 ComputeDiffXyz.
#include "common.cuh"
#include <algorithm>

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// a10 fast path: one warp per (centre, 32-lane chunk of the flattened [m*ofn] axis), 16-byte lanes (4 fp32 / 8 bf16
// columns).  The S = ofn/V consecutive lanes that share an anchor `a` reduce their partial dot products with
// xor-shuffles, replacing the reference's atomicAdd over the ofn threads
// (locationWeightFeatSumBackwardKernel TFNeighborKernel.cu:213-252, atomic at :245).
// Requires ofn a power of two in [V, 32*V] and 16-byte aligned tensors.  Two edges are in flight per step.
template <typename T>
__global__ void __launch_bounds__(kBlock)
k_locw_feat_sum_bwd(const T *__restrict__ lw, const T *__restrict__ tfeats, const T *__restrict__ dsum,
                    const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn, int64_t m, int ofn,
                    int CH, T *__restrict__ dlw, T *__restrict__ dtfeats) {
    constexpr int V = 16 / (int)sizeof(T);
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t i = wid / CH;
    if (i >= pn) return;  // whole warp exits together
    const int ch = (int)(wid - i * CH);
    const int64_t MO = m * ofn;
    const int col = (ch * 32 + lane) * V;
    const bool active = col < MO;
    const int S = ofn / V;
    const int a = active ? col / ofn : 0;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<V> d = vzero<V>();
    if (active) d = IO<T, V>::ldg(dsum + i * MO + col);
    for (int r = 0; r < len; ++r) {
        const int64_t e = bg + r;
        float part = 0.f;
        if (active) {
            const float w = IO<T, 1>::ldg(lw + e * m + a).v[0];
            const Vec<V> f = IO<T, V>::ldg(tfeats + e * MO + col);
            Vec<V> o;
#pragma unroll
            for (int k = 0; k < V; ++k) { o.v[k] = d.v[k] * w; part = fmaf(d.v[k], f.v[k], part); }
            IO<T, V>::stg(dtfeats + e * MO + col, o);
        }
        for (int off = S >> 1; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        if (active && (lane & (S - 1)) == 0) { Vec<1> t; t.v[0] = part; IO<T, 1>::stg(dlw + e * m + a, t); }
    }
}

// a10 generic fallback (any ofn): one thread per (centre, anchor), serial over edges and ofn.
template <typename T>
__global__ void __launch_bounds__(kBlock)
k_locw_feat_sum_bwd_generic(const T *__restrict__ lw, const T *__restrict__ tfeats, const T *__restrict__ dsum,
                            const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn, int64_t m,
                            int64_t ofn, T *__restrict__ dlw, T *__restrict__ dtfeats) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t / m;
    if (i >= pn) return;
    const int64_t a = t - i * m;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const T *dp = dsum + (i * m + a) * ofn;
    for (int r = 0; r < len; ++r) {
        const int64_t e = bg + r;
        const float w = IO<T, 1>::ldg(lw + e * m + a).v[0];
        const T *fp = tfeats + (e * m + a) * ofn;
        T *op = dtfeats + (e * m + a) * ofn;
        float part = 0.f;
        for (int64_t o = 0; o < ofn; ++o) {
            const float dv = IO<T, 1>::ldg(dp + o).v[0];
            Vec<1> ov; ov.v[0] = dv * w;
            IO<T, 1>::stg(op + o, ov);
            part = fmaf(dv, IO<T, 1>::ldg(fp + o).v[0], part);
        }
        Vec<1> pv; pv.v[0] = part;
        IO<T, 1>::stg(dlw + e * m + a, pv);
    }
}

// a11: a sub-warp of G lanes compacts FOUR consecutive rows per step with ballot/popc (order preserved), all four
// rows' index loads in flight together (one row at a time left 6 KB of loads in flight per SM: 26 % of HBM);
// the reference walks the row serially in one thread (eliminateCenter TFNeighborKernelNew.cu:337-370).
constexpr int kElimRows = 4;
__global__ void __launch_bounds__(kBlock)
k_eliminate_center(const int32_t *__restrict__ inidxs, const int32_t *__restrict__ lens,
                   const int32_t *__restrict__ bgs, int64_t pn, int gshift, int32_t *__restrict__ onidxs,
                   int32_t *__restrict__ olens, int32_t *__restrict__ obgs, int32_t *__restrict__ ocidxs) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int G = 1 << gshift;
    const int64_t i0 = (t >> gshift) * kElimRows;           // first row of this group
    const int gl = (int)(t & (G - 1));                      // lane within the group
    const int gbase = (threadIdx.x & 31) & ~(G - 1);        // first warp lane of the group
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << gbase);
    const unsigned below = (1u << (gbase + gl)) - 1u;
    int64_t bg[kElimRows];
    int len[kElimRows], written[kElimRows];
    int maxlen = 0;
#pragma unroll
    for (int k = 0; k < kElimRows; ++k) {
        const bool valid = i0 + k < pn;
        bg[k] = valid ? (int64_t)__ldg(bgs + i0 + k) : 0;
        len[k] = valid ? __ldg(lens + i0 + k) : 0;
        written[k] = 0;
        maxlen = max(maxlen, len[k]);
    }
    // every lane of the warp must reach the ballots: loop to the warp-wide maximum row length
    for (int off = 16; off > 0; off >>= 1) maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, off));
    for (int base = 0; base < maxlen; base += G) {
        const int j = base + gl;
        int32_t v[kElimRows];
#pragma unroll
        for (int k = 0; k < kElimRows; ++k) v[k] = (j < len[k]) ? __ldg(inidxs + bg[k] + j) : -1;
#pragma unroll
        for (int k = 0; k < kElimRows; ++k) {
            const bool keep = (j < len[k]) && (v[k] != (int32_t)(i0 + k));
            const unsigned bal = __ballot_sync(0xffffffffu, keep) & gmask;
            const int pos = written[k] + __popc(bal & below);
            const int64_t obg = bg[k] - (i0 + k);
            if (keep && pos < len[k] - 1) { onidxs[obg + pos] = v[k]; ocidxs[obg + pos] = (int32_t)(i0 + k); }
            written[k] += __popc(bal);
        }
    }
    if (gl < kElimRows && i0 + gl < pn) {
        int l = len[0]; int64_t b = bg[0];
#pragma unroll
        for (int k = 1; k < kElimRows; ++k) { l = (gl == k) ? len[k] : l; b = (gl == k) ? bg[k] : b; }
        olens[i0 + gl] = l - 1;
        obgs[i0 + gl] = (int32_t)(b - (i0 + gl));
    }
}

// a11, edge-parallel: a CTA owns a tile of consecutive centres = (for a gap-free CSR) a contiguous slice of the edge
// array.  Per chunk of 2048 edges: the tile's rows are expanded into a per-edge local row id in shared memory, every
// thread loads 8 edges (coalesced), the self edge of each row is located (pass A), and every kept edge is written at
// e - i - (j > selfpos_i) (pass B): coalesced reads AND writes, no idle lanes whatever the degree.  This is exact for
// rows with at most one self edge (a row without one drops its last entry, as the reference's `pos < len - 1` guard
// does, TFNeighborKernelNew.cu:337-370); rows with several self edges are redone serially by one thread each, and a
// tile whose CSR has gaps goes through the same serial path.
constexpr int kEcT = 256, kEcEPT = 8, kEcChunk = kEcT * kEcEPT, kEcMaxRows = 1024;
__global__ void __launch_bounds__(kEcT)
k_eliminate_center_tile(const int32_t *__restrict__ inidxs, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
                        int64_t pn, int Tc, int32_t *__restrict__ onidxs, int32_t *__restrict__ olens,
                        int32_t *__restrict__ obgs, int32_t *__restrict__ ocidxs) {
    __shared__ int s_bg[kEcMaxRows + 1], s_len[kEcMaxRows], s_self[kEcMaxRows], s_nself[kEcMaxRows];
    __shared__ unsigned short s_row[kEcChunk];
    const int tid = threadIdx.x;
    const int64_t i0 = (int64_t)blockIdx.x * Tc;
    const int n_c = (int)min((int64_t)Tc, pn - i0);
    for (int t = tid; t < n_c; t += kEcT) {
        s_bg[t] = bgs[i0 + t]; s_len[t] = lens[i0 + t];
        s_self[t] = 0x7fffffff; s_nself[t] = 0;
    }
    __syncthreads();
    if (tid == 0) s_bg[n_c] = s_bg[n_c - 1] + s_len[n_c - 1];
    __syncthreads();
    bool ok = true;
    for (int t = tid; t < n_c; t += kEcT) ok = ok && s_len[t] >= 0 && (s_bg[t] + s_len[t] == s_bg[t + 1]);
    ok = __syncthreads_and(ok);
    for (int t = tid; t < n_c; t += kEcT) {  // the per-row outputs
        olens[i0 + t] = s_len[t] - 1;
        obgs[i0 + t] = (int32_t)(s_bg[t] - (i0 + t));
    }
    auto serial_row = [&](int t) {  // generic compaction of one row by one thread (order preserved)
        const int64_t gi = i0 + t;
        const int64_t ob = (int64_t)s_bg[t] - gi;
        int pos = 0;
        for (int j = 0; j < s_len[t]; ++j) {
            const int32_t v = __ldg(inidxs + s_bg[t] + j);
            if (v != (int32_t)gi) {
                if (pos < s_len[t] - 1) { onidxs[ob + pos] = v; ocidxs[ob + pos] = (int32_t)gi; }
                ++pos;
            }
        }
    };
    if (!ok) {
        for (int t = tid; t < n_c; t += kEcT) serial_row(t);
        return;
    }
    // 32-bit arithmetic from here on (csum < 2^31).  s_self holds the ABSOLUTE edge index of a row's first self edge.
    const int e_lo = s_bg[0], e_hi = s_bg[n_c];
    const int gi0 = (int)i0;
    for (int c0 = e_lo; c0 < e_hi; c0 += kEcChunk) {
        const int c1 = min(c0 + kEcChunk, e_hi);
        for (int t = tid; t < n_c; t += kEcT) {  // expand: local row id of every edge of the chunk
            const int lo = max(s_bg[t], c0) - c0, hi = min(s_bg[t + 1], c1) - c0;
            for (int x = lo; x < hi; ++x) s_row[x] = (unsigned short)t;
        }
        __syncthreads();
        int32_t v[kEcEPT];
        int row[kEcEPT];
#pragma unroll
        for (int k = 0; k < kEcEPT; ++k) {
            const int x = tid + k * kEcT;
            const bool in = c0 + x < c1;
            v[k] = in ? __ldg(inidxs + c0 + x) : -1;
            row[k] = in ? (int)s_row[x] : 0;
        }
#pragma unroll
        for (int k = 0; k < kEcEPT; ++k)  // pass A: where is the self edge of each row
            if (v[k] == gi0 + row[k]) { atomicMin(&s_self[row[k]], c0 + tid + k * kEcT); atomicAdd(&s_nself[row[k]], 1); }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < kEcEPT; ++k) {  // pass B: a kept edge moves up by (centres before it) + (self edge passed)
            const int e = c0 + tid + k * kEcT;
            const int gi = gi0 + row[k];
            const int w = e - gi - (e > s_self[row[k]] ? 1 : 0);
            // v = -1 marks an out-of-range slot; w < next row's start - (gi + 1) <=> position < len - 1
            if (v[k] >= 0 && v[k] != gi && w < s_bg[row[k] + 1] - gi - 1) {
                onidxs[w] = v[k];
                ocidxs[w] = gi;
            }
        }
        __syncthreads();  // s_row is rewritten by the next chunk
    }
    for (int t = tid; t < n_c; t += kEcT)
        if (s_nself[t] > 1) serial_row(t);  // several self edges in one row: redo it exactly
}

// cidxs[bg_i + j] = i: 8 lanes per centre.
__global__ void __launch_bounds__(kBlock)
k_csr_cidxs(const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs, int64_t pn, int64_t csum,
            int32_t *__restrict__ cidxs) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t >> 3;
    if (i >= pn) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    for (int j = (int)(t & 7); j < len; j += 8)
        if (bg + j < csum) cidxs[bg + j] = (int32_t)i;
}

// a12: IEEE divide (no fast-math), floor, saturating float->uint32 conversion — the SASS the
// reference compiles to (F2I.U32.FLOOR; computeVoxelIdxKernel PoolingIndex/ComputeVoxelIdx.cu:4-27).
__device__ __forceinline__ int32_t voxel_of(float x, float mn, float voxel_len) {
    return (int32_t)__float2uint_rd(__fdiv_rn(__fsub_rn(x, mn), voxel_len));
}
__global__ void __launch_bounds__(kBlock)
k_voxel_index(const float *__restrict__ pts, int64_t pn, int64_t stride, float voxel_len, float min_x,
              float min_y, float min_z, int32_t *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pn) return;
    const float x = pts[i * stride], y = pts[i * stride + 1], z = pts[i * stride + 2];
    out[i * 3 + 0] = voxel_of(x, min_x, voxel_len);
    out[i * 3 + 1] = voxel_of(y, min_y, voxel_len);
    out[i * 3 + 2] = voxel_of(z, min_z, voxel_len);
}
// packed [pn,3] rows, 16-byte aligned: a thread converts 4 points = three 16-byte loads, three 16-byte stores
__global__ void __launch_bounds__(kBlock)
k_voxel_index4(const float4 *__restrict__ pts, int64_t n_quads, float voxel_len, float min_x, float min_y, float min_z,
               int4 *__restrict__ out) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_quads) return;
    const float4 a = __ldg(pts + q * 3), b = __ldg(pts + q * 3 + 1), c = __ldg(pts + q * 3 + 2);
    // a = x0 y0 z0 x1 | b = y1 z1 x2 y2 | c = z2 x3 y3 z3
    __stcs(out + q * 3, make_int4(voxel_of(a.x, min_x, voxel_len), voxel_of(a.y, min_y, voxel_len),
                                  voxel_of(a.z, min_z, voxel_len), voxel_of(a.w, min_x, voxel_len)));
    __stcs(out + q * 3 + 1, make_int4(voxel_of(b.x, min_y, voxel_len), voxel_of(b.y, min_z, voxel_len),
                                      voxel_of(b.z, min_x, voxel_len), voxel_of(b.w, min_y, voxel_len)));
    __stcs(out + q * 3 + 2, make_int4(voxel_of(c.x, min_z, voxel_len), voxel_of(c.y, min_x, voxel_len),
                                      voxel_of(c.z, min_y, voxel_len), voxel_of(c.w, min_z, voxel_len)));
}

// a16: a single fp32 subtract per coordinate, bit-exact (computeDiffXYZKernel PoolingIndex/ComputeDiffXYZ.cu:4-19).
__global__ void __launch_bounds__(kBlock)
k_diff_xyz(const float *__restrict__ xyzs, const float *__restrict__ cxyzs, const int32_t *__restrict__ cidxs,
           int64_t p0, int64_t pn1, float *__restrict__ out) {
    const int64_t t = p0 * 3 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= pn1 * 3) return;
    const int64_t i = t / 3;
    const int k = (int)(t - i * 3);
    out[t] = __fsub_rn(xyzs[t], __ldg(cxyzs + (int64_t)cidxs[i] * 3 + k));
}
// 4 points per thread: 16-byte loads of the coordinates and of the 4 centre ids, 12 centre gathers in flight
__global__ void __launch_bounds__(kBlock)
k_diff_xyz4(const float4 *__restrict__ xyzs, const float *__restrict__ cxyzs, const int4 *__restrict__ cidxs,
            int64_t n_quads, float4 *__restrict__ out) {
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_quads) return;
    const int4 ci = __ldg(cidxs + q);
    const float4 a = __ldg(xyzs + q * 3), b = __ldg(xyzs + q * 3 + 1), c = __ldg(xyzs + q * 3 + 2);
    const float *c0 = cxyzs + (int64_t)ci.x * 3, *c1 = cxyzs + (int64_t)ci.y * 3, *c2 = cxyzs + (int64_t)ci.z * 3,
                *c3 = cxyzs + (int64_t)ci.w * 3;
    const float x0 = __ldg(c0), y0 = __ldg(c0 + 1), z0 = __ldg(c0 + 2), x1 = __ldg(c1), y1 = __ldg(c1 + 1), z1 = __ldg(c1 + 2);
    const float x2 = __ldg(c2), y2 = __ldg(c2 + 1), z2 = __ldg(c2 + 2), x3 = __ldg(c3), y3 = __ldg(c3 + 1), z3 = __ldg(c3 + 2);
    __stcs(out + q * 3, make_float4(__fsub_rn(a.x, x0), __fsub_rn(a.y, y0), __fsub_rn(a.z, z0), __fsub_rn(a.w, x1)));
    __stcs(out + q * 3 + 1, make_float4(__fsub_rn(b.x, y1), __fsub_rn(b.y, z1), __fsub_rn(b.z, x2), __fsub_rn(b.w, y2)));
    __stcs(out + q * 3 + 2, make_float4(__fsub_rn(c.x, z2), __fsub_rn(c.y, x3), __fsub_rn(c.z, y3), __fsub_rn(c.w, z3)));
}
}  // namespace

static int locw_feat_sum_backward_impl(bool bf, const void *lw, const void *tfeats, const void *dsum,
                                       const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t m, int64_t ofn,
                                       int64_t csum, void *dlw, void *dtfeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum_backward: negative size");
    if (pn == 0 || m == 0 || csum == 0) return 0;
    cudaStream_t s = as_stream(stream);
    const int es = bf ? 2 : 4;
    if (ofn == 0) { PCS_CUDA(cudaMemsetAsync(dlw, 0, (size_t)es * csum * m, s)); return 0; }
    const int V = 16 / es;
    const bool pow2 = (ofn & (ofn - 1)) == 0;
    const bool fast = pow2 && ofn >= V && ofn <= 32 * V && aligned(tfeats, 16) && aligned(dsum, 16) && aligned(dtfeats, 16);
    if (fast) {
        const int CH = (int)((m * ofn + 32 * V - 1) / (32 * V));
        const int64_t warps = pn * CH;
        const int64_t grid = (warps * 32 + kBlock - 1) / kBlock;
        if (bf) k_locw_feat_sum_bwd<bf16><<<grid, kBlock, 0, s>>>((const bf16 *)lw, (const bf16 *)tfeats, (const bf16 *)dsum, lens, bgs, pn, m, (int)ofn, CH, (bf16 *)dlw, (bf16 *)dtfeats);
        else k_locw_feat_sum_bwd<float><<<grid, kBlock, 0, s>>>((const float *)lw, (const float *)tfeats, (const float *)dsum, lens, bgs, pn, m, (int)ofn, CH, (float *)dlw, (float *)dtfeats);
    } else {
        const int64_t grid = (pn * m + kBlock - 1) / kBlock;
        if (bf) k_locw_feat_sum_bwd_generic<bf16><<<grid, kBlock, 0, s>>>((const bf16 *)lw, (const bf16 *)tfeats, (const bf16 *)dsum, lens, bgs, pn, m, ofn, (bf16 *)dlw, (bf16 *)dtfeats);
        else k_locw_feat_sum_bwd_generic<float><<<grid, kBlock, 0, s>>>((const float *)lw, (const float *)tfeats, (const float *)dsum, lens, bgs, pn, m, ofn, (float *)dlw, (float *)dtfeats);
    }
    return check_launch("location_weight_feat_sum_backward");
}
extern "C" int pcs_location_weight_feat_sum_backward(const float *lw, const float *tfeats, const float *dsum,
                                                     const int32_t *lens, const int32_t *bgs, int64_t pn,
                                                     int64_t m, int64_t ofn, int64_t csum, float *dlw,
                                                     float *dtfeats, pcs_stream_t stream) {
    return locw_feat_sum_backward_impl(false, lw, tfeats, dsum, lens, bgs, pn, m, ofn, csum, dlw, dtfeats, stream);
}
extern "C" int pcs_location_weight_feat_sum_backward_bf16(const void *lw, const void *tfeats, const void *dsum,
                                                          const int32_t *lens, const int32_t *bgs, int64_t pn,
                                                          int64_t m, int64_t ofn, int64_t csum, void *dlw,
                                                          void *dtfeats, pcs_stream_t stream) {
    return locw_feat_sum_backward_impl(true, lw, tfeats, dsum, lens, bgs, pn, m, ofn, csum, dlw, dtfeats, stream);
}

extern "C" int pcs_csr_cidxs(const int32_t *lens, const int32_t *bgs, int64_t pn, int64_t csum, int32_t *cidxs,
                             pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && csum >= 0, "csr_cidxs: negative size");
    if (pn == 0 || csum == 0) return 0;
    k_csr_cidxs<<<(unsigned)((pn * 8 + kBlock - 1) / kBlock), kBlock, 0, as_stream(stream)>>>(lens, bgs, pn, csum, cidxs);
    return check_launch("csr_cidxs");
}

extern "C" int pcs_eliminate_center(const int32_t *inidxs, const int32_t *lens, const int32_t *bgs, int64_t pn,
                                    int64_t csum, int32_t *onidxs, int32_t *olens, int32_t *obgs,
                                    int32_t *ocidxs, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && csum >= 0, "eliminate_center: negative size");
    if (pn == 0) return 0;
    const int64_t avg = (csum + pn - 1) / pn;
    if (csum < ((int64_t)1 << 31) && opt::elim_tile.load()) {
        // edge-parallel tile kernel: ~two chunks of edges per tile (measured: 1 chunk 62.8 %, 2 chunks 66.0 % of HBM)
        const int Tc = (int)std::max<int64_t>(1, std::min<int64_t>(kEcMaxRows, 2 * kEcChunk / std::max<int64_t>(avg, 1)));
        const int64_t tiles = (pn + Tc - 1) / Tc;
        PCS_REQUIRE(tiles < ((int64_t)1 << 31), "eliminate_center: too many tiles");
        k_eliminate_center_tile<<<(unsigned)tiles, kEcT, 0, as_stream(stream)>>>(inidxs, lens, bgs, pn, Tc, onidxs, olens, obgs, ocidxs);
        return check_launch("eliminate_center");
    }
    int G = avg <= 8 ? 8 : (avg <= 16 ? 16 : 32);
    const int gshift = ilog2(G);
    const int64_t groups = (pn + kElimRows - 1) / kElimRows;
    const int64_t grid = ((groups << gshift) + kBlock - 1) / kBlock;
    k_eliminate_center<<<grid, kBlock, 0, as_stream(stream)>>>(inidxs, lens, bgs, pn, gshift, onidxs, olens, obgs, ocidxs);
    return check_launch("eliminate_center");
}

extern "C" int pcs_compute_voxel_index(const float *pts, int64_t pn, int64_t pt_stride, float voxel_len,
                                       float min_x, float min_y, float min_z, int32_t *voxel_idxs,
                                       pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pt_stride >= 3, "compute_voxel_index: bad size");
    if (pn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    int64_t done = 0;
    if (pt_stride == 3 && aligned(pts, 16) && aligned(voxel_idxs, 16) && pn >= 4) {
        const int64_t nq = pn / 4;
        k_voxel_index4<<<(nq + kBlock - 1) / kBlock, kBlock, 0, s>>>(reinterpret_cast<const float4 *>(pts), nq, voxel_len, min_x, min_y,
                                                                 min_z, reinterpret_cast<int4 *>(voxel_idxs));
        if (check_launch("compute_voxel_index")) return 1;
        done = nq * 4;
    }
    if (done < pn) {
        const int64_t rest = pn - done;
        k_voxel_index<<<(rest + kBlock - 1) / kBlock, kBlock, 0, s>>>(pts + done * pt_stride, rest, pt_stride, voxel_len, min_x, min_y,
                                                                  min_z, voxel_idxs + done * 3);
        return check_launch("compute_voxel_index");
    }
    return 0;
}

extern "C" int pcs_compute_diff_xyz(const float *xyzs, const float *cxyzs, const int32_t *cidxs, int64_t pn1,
                                    int64_t pn2, float *dxyzs, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0, "compute_diff_xyz: negative size");
    if (pn1 == 0) return 0;
    cudaStream_t s = as_stream(stream);
    int64_t done = 0;
    if (aligned(xyzs, 16) && aligned(dxyzs, 16) && aligned(cidxs, 16) && pn1 >= 4) {
        const int64_t nq = pn1 / 4;
        k_diff_xyz4<<<(nq + kBlock - 1) / kBlock, kBlock, 0, s>>>(reinterpret_cast<const float4 *>(xyzs), cxyzs,
                                                              reinterpret_cast<const int4 *>(cidxs), nq, reinterpret_cast<float4 *>(dxyzs));
        if (check_launch("compute_diff_xyz")) return 1;
        done = nq * 4;
    }
    if (done < pn1) {
        const int64_t rest = (pn1 - done) * 3;
        k_diff_xyz<<<(rest + kBlock - 1) / kBlock, kBlock, 0, s>>>(xyzs, cxyzs, cidxs, done, pn1, dxyzs);
        return check_launch("compute_diff_xyz");
    }
    return 0;
}

// neighbor_misc.cu — a10 LocationWeightFeatSumBackward, a11 EliminateCenter, a12 ComputeVoxelIndex,
// a16 ComputeDiffXyz.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// a10 fast path: one warp per (centre, 128-column chunk of the flattened [m*ofn] axis), float4
// lanes.  The S = ofn/4 consecutive lanes that share an anchor `a` reduce their partial dot
// products with xor-shuffles, replacing the reference's atomicAdd over the ofn threads
// (locationWeightFeatSumBackwardKernel TFNeighborKernel.cu:213-252, atomic at :245).
// Requires ofn a power of two in [4,128] and 16-byte aligned tensors.
__global__ void __launch_bounds__(kBlock)
k_locw_feat_sum_bwd(const float *__restrict__ lw, const float *__restrict__ tfeats,
                    const float *__restrict__ dsum, const int32_t *__restrict__ lens,
                    const int32_t *__restrict__ bgs, int64_t pn, int64_t m, int ofn, int CH,
                    float *__restrict__ dlw, float *__restrict__ dtfeats) {
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const int64_t i = wid / CH;
    if (i >= pn) return;  // whole warp exits together
    const int ch = (int)(wid - i * CH);
    const int64_t MO = m * ofn;
    const int col = (ch * 32 + lane) * 4;
    const bool active = col < MO;
    const int S = ofn >> 2;
    const int a = active ? col / ofn : 0;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    Vec<4> d = vzero<4>();
    if (active) d = ld<4>(dsum + i * MO + col);
    for (int r = 0; r < len; ++r) {
        const int64_t e = bg + r;
        float part = 0.f;
        if (active) {
            const float w = __ldg(lw + e * m + a);
            const Vec<4> f = ld<4>(tfeats + e * MO + col);
            Vec<4> o;
#pragma unroll
            for (int k = 0; k < 4; ++k) { o.v[k] = d.v[k] * w; part = fmaf(d.v[k], f.v[k], part); }
            st<4>(dtfeats + e * MO + col, o);
        }
        for (int off = S >> 1; off > 0; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        if (active && (lane & (S - 1)) == 0) dlw[e * m + a] = part;
    }
}

// a10 generic fallback (any ofn): one thread per (centre, anchor), serial over edges and ofn.
__global__ void __launch_bounds__(kBlock)
k_locw_feat_sum_bwd_generic(const float *__restrict__ lw, const float *__restrict__ tfeats,
                            const float *__restrict__ dsum, const int32_t *__restrict__ lens,
                            const int32_t *__restrict__ bgs, int64_t pn, int64_t m, int64_t ofn,
                            float *__restrict__ dlw, float *__restrict__ dtfeats) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t / m;
    if (i >= pn) return;
    const int64_t a = t - i * m;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const float *dp = dsum + (i * m + a) * ofn;
    for (int r = 0; r < len; ++r) {
        const int64_t e = bg + r;
        const float w = lw[e * m + a];
        const float *fp = tfeats + (e * m + a) * ofn;
        float *op = dtfeats + (e * m + a) * ofn;
        float part = 0.f;
        for (int64_t o = 0; o < ofn; ++o) {
            const float dv = dp[o];
            op[o] = dv * w;
            part = fmaf(dv, fp[o], part);
        }
        dlw[e * m + a] = part;
    }
}

// a11: a sub-warp of G lanes compacts one row with ballot/popc (order preserved); the reference
// walks the row serially in one thread (eliminateCenter TFNeighborKernelNew.cu:337-370).
__global__ void __launch_bounds__(kBlock)
k_eliminate_center(const int32_t *__restrict__ inidxs, const int32_t *__restrict__ lens,
                   const int32_t *__restrict__ bgs, int64_t pn, int gshift, int32_t *__restrict__ onidxs,
                   int32_t *__restrict__ olens, int32_t *__restrict__ obgs, int32_t *__restrict__ ocidxs) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int G = 1 << gshift;
    const int64_t i = t >> gshift;
    const int gl = (int)(t & (G - 1));                      // lane within the group
    const int gbase = (threadIdx.x & 31) & ~(G - 1);        // first warp lane of the group
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << gbase);
    const bool valid = i < pn;
    const int64_t bg = valid ? (int64_t)bgs[i] : 0;
    const int len = valid ? lens[i] : 0;
    const int64_t obg = bg - i;
    const int cap = len - 1;
    // every lane of the warp must reach the ballots: loop to the warp-wide maximum row length
    int maxlen = len;
    for (int off = 16; off > 0; off >>= 1) maxlen = max(maxlen, __shfl_xor_sync(0xffffffffu, maxlen, off));
    int written = 0;
    for (int base = 0; base < maxlen; base += G) {
        const int j = base + gl;
        int32_t v = 0;
        bool keep = false;
        if (j < len) { v = inidxs[bg + j]; keep = (v != (int32_t)i); }
        const unsigned bal = __ballot_sync(0xffffffffu, keep) & gmask;
        const int pos = written + __popc(bal & ((1u << (gbase + gl)) - 1u));
        if (keep && pos < cap) { onidxs[obg + pos] = v; ocidxs[obg + pos] = (int32_t)i; }
        written += __popc(bal);
    }
    if (valid && gl == 0) { olens[i] = len - 1; obgs[i] = (int32_t)obg; }
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
__global__ void __launch_bounds__(kBlock)
k_voxel_index(const float *__restrict__ pts, int64_t pn, int64_t stride, float voxel_len, float min_x,
              float min_y, float min_z, int32_t *__restrict__ out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pn) return;
    const float x = pts[i * stride], y = pts[i * stride + 1], z = pts[i * stride + 2];
    out[i * 3 + 0] = (int32_t)__float2uint_rd(__fdiv_rn(__fsub_rn(x, min_x), voxel_len));
    out[i * 3 + 1] = (int32_t)__float2uint_rd(__fdiv_rn(__fsub_rn(y, min_y), voxel_len));
    out[i * 3 + 2] = (int32_t)__float2uint_rd(__fdiv_rn(__fsub_rn(z, min_z), voxel_len));
}

// a16: one thread per coordinate (3 per point) — a single fp32 subtract, bit-exact
// (computeDiffXYZKernel PoolingIndex/ComputeDiffXYZ.cu:4-19).
__global__ void __launch_bounds__(kBlock)
k_diff_xyz(const float *__restrict__ xyzs, const float *__restrict__ cxyzs, const int32_t *__restrict__ cidxs,
           int64_t pn1, float *__restrict__ out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= pn1 * 3) return;
    const int64_t i = t / 3;
    const int k = (int)(t - i * 3);
    out[t] = __fsub_rn(xyzs[t], __ldg(cxyzs + (int64_t)cidxs[i] * 3 + k));
}
}  // namespace

extern "C" int pcs_location_weight_feat_sum_backward(const float *lw, const float *tfeats, const float *dsum,
                                                     const int32_t *lens, const int32_t *bgs, int64_t pn,
                                                     int64_t m, int64_t ofn, int64_t csum, float *dlw,
                                                     float *dtfeats, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && m >= 0 && ofn >= 0 && csum >= 0, "location_weight_feat_sum_backward: negative size");
    if (pn == 0 || m == 0 || csum == 0) return 0;
    cudaStream_t s = as_stream(stream);
    if (ofn == 0) { PCS_CUDA(cudaMemsetAsync(dlw, 0, sizeof(float) * csum * m, s)); return 0; }
    const bool pow2 = (ofn & (ofn - 1)) == 0;
    const bool fast = pow2 && ofn >= 4 && ofn <= 128 && aligned(tfeats, 16) && aligned(dsum, 16) && aligned(dtfeats, 16);
    if (fast) {
        const int CH = (int)((m * ofn + 127) / 128);
        const int64_t warps = pn * CH;
        const int64_t grid = (warps * 32 + kBlock - 1) / kBlock;
        k_locw_feat_sum_bwd<<<grid, kBlock, 0, s>>>(lw, tfeats, dsum, lens, bgs, pn, m, (int)ofn, CH, dlw, dtfeats);
    } else {
        const int64_t grid = (pn * m + kBlock - 1) / kBlock;
        k_locw_feat_sum_bwd_generic<<<grid, kBlock, 0, s>>>(lw, tfeats, dsum, lens, bgs, pn, m, ofn, dlw, dtfeats);
    }
    return check_launch("location_weight_feat_sum_backward");
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
    int G = avg <= 8 ? 8 : (avg <= 16 ? 16 : 32);
    const int gshift = ilog2(G);
    const int64_t grid = ((pn << gshift) + kBlock - 1) / kBlock;
    k_eliminate_center<<<grid, kBlock, 0, as_stream(stream)>>>(inidxs, lens, bgs, pn, gshift, onidxs, olens, obgs, ocidxs);
    return check_launch("eliminate_center");
}

extern "C" int pcs_compute_voxel_index(const float *pts, int64_t pn, int64_t pt_stride, float voxel_len,
                                       float min_x, float min_y, float min_z, int32_t *voxel_idxs,
                                       pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pt_stride >= 3, "compute_voxel_index: bad size");
    if (pn == 0) return 0;
    k_voxel_index<<<(pn + kBlock - 1) / kBlock, kBlock, 0, as_stream(stream)>>>(pts, pn, pt_stride, voxel_len, min_x, min_y, min_z, voxel_idxs);
    return check_launch("compute_voxel_index");
}

extern "C" int pcs_compute_diff_xyz(const float *xyzs, const float *cxyzs, const int32_t *cidxs, int64_t pn1,
                                    int64_t pn2, float *dxyzs, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0, "compute_diff_xyz: negative size");
    if (pn1 == 0) return 0;
    k_diff_xyz<<<(pn1 * 3 + kBlock - 1) / kBlock, kBlock, 0, as_stream(stream)>>>(xyzs, cxyzs, cidxs, pn1, dxyzs);
    return check_launch("compute_diff_xyz");
}

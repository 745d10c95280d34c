// point_util.cu — the libPointUtil calls the reference makes either side of the hot path:
// block preparation (s3dis_util.py:33-89: computeCovarsGPU; aug_util.py:85-101: gatherBlockPointsGPU) and label
// interpolation (interpolate.py:135-142: interpolateProbsGPU).  libPointUtil is an un-vendored dependency with no
// pinned version: the semantics here are restated from the call sites (argument lists, shapes, how the results are
// used) and stated in include/pcseg_ops.h; parity is pinned by the numpy oracle only ("parity unpinned by the
// reference").  All three are HBM-bound byte movers over CSR neighbour lists, built like the a-row kernels.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// ---- covariance features: 8 lanes per query walk its neighbour list twice (mean, then centred outer products);
// out[q] = the 3x3 covariance (row-major, 9 floats) scaled to unit Frobenius norm (zeros for < 2 neighbours or a
// degenerate neighbourhood).  The second pass re-reads rows that the first one left in L1/L2.
__global__ void __launch_bounds__(kBlock)
k_compute_covars(const float *__restrict__ xyzs, const int32_t *__restrict__ nidxs, const int32_t *__restrict__ lens,
                 const int32_t *__restrict__ bgs, int64_t qn, float *__restrict__ covars) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t q = t >> 3;
    const int gl = (int)(t & 7);
    const bool valid = q < qn;  // whole 8-lane groups are valid or not; shuffles below stay within the group
    const int len = valid ? __ldg(lens + q) : 0;
    const int32_t *np = nidxs + (valid ? (int64_t)__ldg(bgs + q) : 0);
    float sx = 0.f, sy = 0.f, sz = 0.f;
    for (int r = gl; r < len; r += 8) {
        const int64_t n = __ldg(np + r);
        sx += __ldg(xyzs + n * 3); sy += __ldg(xyzs + n * 3 + 1); sz += __ldg(xyzs + n * 3 + 2);
    }
#pragma unroll
    for (int off = 4; off > 0; off >>= 1) {
        sx += __shfl_xor_sync(0xffffffffu, sx, off); sy += __shfl_xor_sync(0xffffffffu, sy, off); sz += __shfl_xor_sync(0xffffffffu, sz, off);
    }
    const float inv = len > 0 ? 1.f / (float)len : 0.f;
    const float mx = sx * inv, my = sy * inv, mz = sz * inv;
    float c[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // xx xy xz yy yz zz
    for (int r = gl; r < len; r += 8) {
        const int64_t n = __ldg(np + r);
        const float dx = __ldg(xyzs + n * 3) - mx, dy = __ldg(xyzs + n * 3 + 1) - my, dz = __ldg(xyzs + n * 3 + 2) - mz;
        c[0] += dx * dx; c[1] += dx * dy; c[2] += dx * dz; c[3] += dy * dy; c[4] += dy * dz; c[5] += dz * dz;
    }
#pragma unroll
    for (int k = 0; k < 6; ++k)
#pragma unroll
        for (int off = 4; off > 0; off >>= 1) c[k] += __shfl_xor_sync(0xffffffffu, c[k], off);
    if (!valid) return;
    const float fro2 = c[0] * c[0] + c[3] * c[3] + c[5] * c[5] + 2.f * (c[1] * c[1] + c[2] * c[2] + c[4] * c[4]);
    const float s = fro2 > 0.f ? rsqrtf(fro2) : 0.f;
    // lane gl writes element gl (and lane 0 also element 8): [xx xy xz | xy yy yz | xz yz zz]
    const float m9[9] = {c[0], c[1], c[2], c[1], c[3], c[4], c[2], c[4], c[5]};
    float v = m9[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) v = gl == k ? m9[k] : v;
    covars[q * 9 + gl] = v * s;
    if (gl == 0) covars[q * 9 + 8] = m9[8] * s;
}

// ---- label interpolation: qprobs[q, :] = sum_k w_k * sprobs[n_k, :] / sum_k w_k,  w_k = exp(-|q - s_{n_k}|^2 * ratio).
// A group of G lanes (G = pow2 >= classes, <= 32) owns a query: lane = class column (coalesced rows of sprobs),
// the weight of each neighbour is computed once per lane (same value in every lane of the group).
__global__ void __launch_bounds__(kBlock)
k_interpolate_probs(const float *__restrict__ sxyzs, const float *__restrict__ qxyzs, const float *__restrict__ sprobs,
                    const int32_t *__restrict__ nidxs, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
                    int64_t qn, int C, int gshift, float ratio, float *__restrict__ qprobs) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t q = t >> gshift;
    const int lane = (int)(t & ((1 << gshift) - 1));
    if (q >= qn) return;
    const float qx = __ldg(qxyzs + q * 3), qy = __ldg(qxyzs + q * 3 + 1), qz = __ldg(qxyzs + q * 3 + 2);
    const int len = __ldg(lens + q);
    const int32_t *np = nidxs + (int64_t)__ldg(bgs + q);
    for (int col = lane; col < C; col += (1 << gshift)) {
        float acc = 0.f, wsum = 0.f;
        for (int r = 0; r < len; ++r) {
            const int64_t n = __ldg(np + r);
            const float dx = __ldg(sxyzs + n * 3) - qx, dy = __ldg(sxyzs + n * 3 + 1) - qy, dz = __ldg(sxyzs + n * 3 + 2) - qz;
            const float w = expf(-(dx * dx + dy * dy + dz * dz) * ratio);
            acc += w * __ldg(sprobs + n * C + col);
            wsum += w;
        }
        // every weight underflowed (all neighbours many sigmas away): fall back to the plain mean of the rows
        if (!(wsum > 0.f) && len > 0) {
            acc = 0.f;
            for (int r = 0; r < len; ++r) acc += __ldg(sprobs + (int64_t)__ldg(np + r) * C + col);
            wsum = (float)len;
        }
        qprobs[q * C + col] = len > 0 ? acc / wsum : 0.f;
    }
}

// ---- block cropping: membership of every point in every block [bx, bx+size) x [by, by+size) --------------------
// Pass 1 counts, per (block, 256-point tile), the member points; the host scans the small [nb, n_tiles] table;
// pass 2 writes the member indices in ascending point order (= np.nonzero, aug_util.py:75-80).
template <bool FILL>
__global__ void __launch_bounds__(kBlock)
k_block_members(const float *__restrict__ xyzs, int64_t pn, const float *__restrict__ begs, int nb, float size,
                int64_t n_tiles, int32_t *__restrict__ tile_counts, const int64_t *__restrict__ tile_offs,
                int32_t *__restrict__ idxs) {
    __shared__ int s_warp[kBlock / 32];
    const int64_t tile = blockIdx.x;
    const int64_t i = tile * kBlock + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float x = 0.f, y = 0.f;
    if (i < pn) { x = __ldg(xyzs + i * 3); y = __ldg(xyzs + i * 3 + 1); }
    for (int b = 0; b < nb; ++b) {
        const float bx = __ldg(begs + 2 * b), by = __ldg(begs + 2 * b + 1);
        const bool m = i < pn && x >= bx && x < bx + size && y >= by && y < by + size;
        const unsigned bal = __ballot_sync(0xffffffffu, m);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < kBlock / 32; ++w) { const int c = s_warp[w]; before += w < warp ? c : 0; total += c; }
        if (FILL) {
            if (m) idxs[tile_offs[(int64_t)b * n_tiles + tile] + before + __popc(bal & ((1u << lane) - 1u))] = (int32_t)i;
        } else if (threadIdx.x == 0) {
            tile_counts[(int64_t)b * n_tiles + tile] = total;
        }
        __syncthreads();
    }
}
}  // namespace

extern "C" int pcs_compute_covars(const float *xyzs, const int32_t *nidxs, const int32_t *lens, const int32_t *bgs,
                                  int64_t qn, float *covars, pcs_stream_t stream) {
    PCS_REQUIRE(qn >= 0, "compute_covars: negative size");
    if (qn == 0) return 0;
    k_compute_covars<<<(unsigned)((qn * 8 + kBlock - 1) / kBlock), kBlock, 0, as_stream(stream)>>>(xyzs, nidxs, lens, bgs, qn, covars);
    return check_launch("compute_covars");
}

extern "C" int pcs_interpolate_probs(const float *sxyzs, const float *qxyzs, const float *sprobs, const int32_t *nidxs,
                                     const int32_t *lens, const int32_t *bgs, int64_t qn, int64_t classes, float ratio,
                                     float *qprobs, pcs_stream_t stream) {
    PCS_REQUIRE(qn >= 0 && classes >= 1 && classes < (1 << 20), "interpolate_probs: bad size");
    if (qn == 0) return 0;
    const int G = classes >= 32 ? 32 : next_pow2(classes);
    const int gshift = ilog2(G);
    const int64_t threads = qn << gshift;
    k_interpolate_probs<<<(unsigned)((threads + kBlock - 1) / kBlock), kBlock, 0, as_stream(stream)>>>(
        sxyzs, qxyzs, sprobs, nidxs, lens, bgs, qn, (int)classes, gshift, ratio, qprobs);
    return check_launch("interpolate_probs");
}

extern "C" int64_t pcs_block_members_tiles(int64_t pn) { return (pn + kBlock - 1) / kBlock; }

extern "C" int pcs_block_members_count(const float *xyzs, int64_t pn, const float *block_begs, int64_t nb, float block_size,
                                       int32_t *tile_counts, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && nb >= 0 && nb < (1 << 20), "block_members_count: bad size");
    if (pn == 0 || nb == 0) return 0;
    const int64_t nt = pcs_block_members_tiles(pn);
    k_block_members<false><<<(unsigned)nt, kBlock, 0, as_stream(stream)>>>(xyzs, pn, block_begs, (int)nb, block_size, nt, tile_counts,
                                                                           nullptr, nullptr);
    return check_launch("block_members_count");
}

extern "C" int pcs_block_members_fill(const float *xyzs, int64_t pn, const float *block_begs, int64_t nb, float block_size,
                                      const int64_t *tile_offs, int32_t *idxs, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && nb >= 0 && nb < (1 << 20), "block_members_fill: bad size");
    if (pn == 0 || nb == 0) return 0;
    const int64_t nt = pcs_block_members_tiles(pn);
    k_block_members<true><<<(unsigned)nt, kBlock, 0, as_stream(stream)>>>(xyzs, pn, block_begs, (int)nb, block_size, nt, nullptr,
                                                                          tile_offs, idxs);
    return check_launch("block_members_fill");
}

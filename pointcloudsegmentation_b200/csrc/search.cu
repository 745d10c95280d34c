// search.cu — a18/a19/a20 radius-neighbour construction, tiled brute force.
// A CTA stages 1024 candidates (SoA) in shared memory; each warp owns 8 centres and tests 32
// candidates per step (lane = candidate), so hits come out of __ballot_sync already in ascending
// candidate order: rows are written coalesced, in the reference's order, with no sort.  The
// reference runs one thread per centre over all pn points straight from global memory, twice
// (SearchNeighborhood.cu:6-28,58-88).  Predicate: identical fp32 expression AND contraction
// (checked against the reference's SASS): d2 = fma(dz,dz, fma(dx,dx, dy*dy)), d = candidate-centre.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kWarps = 8;          // warps per CTA
constexpr int kCPW = 8;            // centres per warp
constexpr int kCentres = kWarps * kCPW;
constexpr int kTile = 1024;        // candidates per shared-memory tile
constexpr int kPad = kTile + 11;   // array pitch (bank spread for the AoS->SoA store)

enum { MODE_COUNT = 0, MODE_FILL = 1, MODE_FIXED = 2 };

__device__ __forceinline__ bool hit_test(float tx, float ty, float tz, float cx, float cy, float cz,
                                         float sq_min, float sq_max, bool use_min) {
    const float dx = __fsub_rn(tx, cx), dy = __fsub_rn(ty, cy), dz = __fsub_rn(tz, cz);
    const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    return use_min ? (d2 < sq_max && d2 > sq_min) : (d2 < sq_max);
}

template <int MODE>
__global__ void __launch_bounds__(kWarps * 32)
k_search(const float *__restrict__ xyzs, int64_t pn, const int32_t *__restrict__ block_begs, float sq_min,
         float sq_max, int use_min, int fixed_size, const int32_t *__restrict__ begs,
         int32_t *__restrict__ lens, int32_t *__restrict__ idxs, int32_t *__restrict__ cens) {
    __shared__ float sm[3 * kPad];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t bb = 0, be = pn;
    if (block_begs) { bb = block_begs[blockIdx.y]; be = block_begs[blockIdx.y + 1]; }
    const int64_t c_first = bb + (int64_t)blockIdx.x * kCentres;
    if (c_first >= be) return;  // whole CTA
    const int64_t c0 = c_first + warp * kCPW;

    float cx[kCPW], cy[kCPW], cz[kCPW];
    int cnt[kCPW];
    int64_t wpos[kCPW];
#pragma unroll
    for (int c = 0; c < kCPW; ++c) {
        const int64_t ci = c0 + c;
        const bool v = ci < be;
        cx[c] = v ? __ldg(xyzs + ci * 3) : 0.f;
        cy[c] = v ? __ldg(xyzs + ci * 3 + 1) : 0.f;
        cz[c] = v ? __ldg(xyzs + ci * 3 + 2) : 0.f;
        cnt[c] = 0;
        wpos[c] = 0;
        if (MODE == MODE_FILL && v) wpos[c] = begs[ci];
        if (MODE == MODE_FIXED) wpos[c] = ci * fixed_size;
    }
    const unsigned lt = (1u << lane) - 1u;
    int last[kCPW];  // MODE_FIXED: last hit written
#pragma unroll
    for (int c = 0; c < kCPW; ++c) last[c] = (int)(c0 + c);

    for (int64_t t0 = bb; t0 < be; t0 += kTile) {
        const int nt = (int)min((int64_t)kTile, be - t0);
        __syncthreads();
        for (int f = threadIdx.x; f < nt * 3; f += kWarps * 32) {
            const int pt = f / 3, comp = f - pt * 3;
            sm[comp * kPad + pt] = __ldg(xyzs + t0 * 3 + f);
        }
        __syncthreads();
        if (c0 < be) {
            for (int j0 = 0; j0 < nt; j0 += 32) {
                const int j = j0 + lane;
                const bool jv = j < nt;
                const float tx = sm[j < kTile ? j : 0], ty = sm[kPad + (j < kTile ? j : 0)],
                            tz = sm[2 * kPad + (j < kTile ? j : 0)];
#pragma unroll
                for (int c = 0; c < kCPW; ++c) {
                    const bool h = jv && (c0 + c < be) && hit_test(tx, ty, tz, cx[c], cy[c], cz[c], sq_min, sq_max, use_min);
                    const unsigned bal = __ballot_sync(0xffffffffu, h);
                    if (MODE == MODE_COUNT) {
                        cnt[c] += __popc(bal);
                    } else if (MODE == MODE_FILL) {
                        if (h) {
                            const int64_t w = wpos[c] + __popc(bal & lt);
                            idxs[w] = (int32_t)(t0 + j);
                            cens[w] = (int32_t)(c0 + c);
                        }
                        wpos[c] += __popc(bal);
                    } else {
                        const int slot = cnt[c] + __popc(bal & lt);
                        if (h && slot < fixed_size) {
                            idxs[wpos[c] + slot] = (int32_t)(t0 + j);
                            cens[wpos[c] + slot] = (int32_t)(c0 + c);
                        }
                        // last hit that made it into the row (warp-uniform)
                        const int room = fixed_size - cnt[c];
                        if (bal && room > 0) {
                            unsigned b2 = bal;
                            const int nh = __popc(bal);
                            if (nh > room) {  // keep only the first `room` set bits
                                for (int k = 0; k < nh - room; ++k) b2 &= ~(1u << (31 - __clz(b2)));
                            }
                            last[c] = (int)(t0 + j0 + (31 - __clz(b2)));
                        }
                        cnt[c] += __popc(bal);
                    }
                }
            }
        }
        if (MODE == MODE_FIXED) {
            // early exit once every centre of the CTA has its K hits (reference: loop condition
            // counts<fixed_size, SearchNeighborhoodFixed.cu:26)
            bool done = true;
#pragma unroll
            for (int c = 0; c < kCPW; ++c) done = done && (c0 + c >= be || cnt[c] >= fixed_size);
            if (__syncthreads_and(done)) break;
        }
    }
    if (c0 >= be) return;
    if (MODE == MODE_COUNT) {
#pragma unroll
        for (int c = 0; c < kCPW; ++c)
            if (lane == c && c0 + c < be) lens[c0 + c] = cnt[c];
    }
    if (MODE == MODE_FIXED) {
#pragma unroll
        for (int c = 0; c < kCPW; ++c) {
            if (c0 + c >= be) continue;
            const int have = min(cnt[c], fixed_size);
            for (int s = have + lane; s < fixed_size; s += 32) {
                idxs[wpos[c] + s] = last[c];
                cens[wpos[c] + s] = (int32_t)(c0 + c);
            }
            if (lane == 0) { lens[c0 + c] = fixed_size; }
        }
    }
}

__global__ void k_fixed_begs(int32_t *__restrict__ begs, int64_t pn, int fixed_size) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < pn) begs[i] = (int32_t)(i * fixed_size);
}

dim3 search_grid(int64_t pn, int64_t nb, int64_t max_block_pn, bool batched) {
    const int64_t per = batched ? max_block_pn : pn;
    return dim3((unsigned)((per + kCentres - 1) / kCentres), (unsigned)(batched ? nb : 1), 1);
}
}  // namespace

extern "C" size_t pcs_search_workspace_bytes(int64_t pn) { return pcs_exclusive_scan_workspace_bytes(pn); }

static int check_blocks(const int32_t *block_begs, int64_t nb, int64_t max_block_pn, int64_t pn) {
    if (block_begs) {
        PCS_REQUIRE(nb >= 1 && nb <= 65535, "search: nb must be in [1,65535]");
        PCS_REQUIRE(max_block_pn >= 1 && max_block_pn <= pn, "search: bad max_block_pn");
    }
    return 0;
}

extern "C" int pcs_search_neighborhood_count(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                             int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                             int32_t *lens, int32_t *begs, int32_t *total_dev, void *workspace,
                                             size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_count: bad pn");
    cudaStream_t s = as_stream(stream);
    if (pn == 0) { PCS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), s)); return 0; }
    if (check_blocks(block_begs, nb, max_block_pn, pn)) return 1;
    k_search<MODE_COUNT><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, s>>>(
        xyzs, pn, block_begs, sq_min, sq_max, use_min, 0, nullptr, lens, nullptr, nullptr);
    if (check_launch("search_count")) return 1;
    return pcs_exclusive_scan_i32(lens, pn, begs, total_dev, workspace, workspace_bytes, stream);
}

extern "C" int pcs_search_neighborhood_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                            int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                            const int32_t *begs, int32_t *idxs, int32_t *cens, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_fill: bad pn");
    if (pn == 0) return 0;
    if (check_blocks(block_begs, nb, max_block_pn, pn)) return 1;
    k_search<MODE_FILL><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, as_stream(stream)>>>(
        xyzs, pn, block_begs, sq_min, sq_max, use_min, 0, begs, nullptr, idxs, cens);
    return check_launch("search_fill");
}

extern "C" int pcs_search_neighborhood_fixed(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                             int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                             int64_t fixed_size, int32_t *idxs, int32_t *lens, int32_t *begs,
                                             int32_t *cens, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fixed_size >= 0 && pn * fixed_size < ((int64_t)1 << 31), "search_fixed: bad size");
    if (pn == 0) return 0;
    if (check_blocks(block_begs, nb, max_block_pn, pn)) return 1;
    cudaStream_t s = as_stream(stream);
    k_search<MODE_FIXED><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, s>>>(
        xyzs, pn, block_begs, sq_min, sq_max, use_min, (int)fixed_size, nullptr, lens, idxs, cens);
    if (check_launch("search_fixed")) return 1;
    k_fixed_begs<<<(pn + 255) / 256, 256, 0, s>>>(begs, pn, (int)fixed_size);
    return check_launch("fixed_begs");
}

// search.cu — a18/a19/a20 radius-neighbour construction: brute force with chunk culling.
// Candidates are visited in ascending index order in chunks of 32 (lane = candidate), so hits come
// out of __ballot_sync already in the reference's order: rows are written in ascending index with
// no sort.  A pre-pass stores the bounding box of every 32-point chunk; a warp (8 consecutive
// centres) first tests 32 chunk boxes at a time against its centres' box inflated by r_max*(1+1e-3)
// and only loads / distance-tests the chunks that overlap.  The cull is conservative (a skipped
// chunk lies > r_max away on one axis for every centre of the warp, so the exact fp32 predicate
// would be false), hence the output is bit-identical to plain brute force for ANY point order; when
// the points are spatially coherent in index order — which they are after the voxel permutation of
// points_pooling — ~90 % of the pn^2 tests disappear.  The boxes are two-level (32-point chunks, and
// super-chunks of 32 chunks = 1024 points): a warp tests 32 super boxes per ballot and descends only into
// the overlapping ones, so the cull itself is O(pn/1024) per centre group instead of O(pn/32) — what
// keeps a 1M-point Semantic3D block at a few ms per search.  The reference runs one thread per centre over
// all pn points straight from global memory, twice (SearchNeighborhood.cu:6-28,58-88).
// Predicate: identical fp32 expression AND contraction (checked against the reference's SASS):
// d2 = fma(dz,dz, fma(dx,dx, dy*dy)), d = candidate - centre.
#include "common.cuh"
#include "search_boxes.cuh"

using namespace pcs;

namespace {
constexpr int kWarps = 8;          // warps per CTA
constexpr int kCPW = 8;            // centres per warp
constexpr int kCentres = kWarps * kCPW;

enum { MODE_COUNT = 0, MODE_FILL = 1, MODE_FIXED = 2 };

__device__ __forceinline__ bool hit_test(float tx, float ty, float tz, float cx, float cy, float cz,
                                         float sq_min, float sq_max, bool use_min) {
    const float dx = __fsub_rn(tx, cx), dy = __fsub_rn(ty, cy), dz = __fsub_rn(tz, cz);
    const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    return use_min ? (d2 < sq_max && d2 > sq_min) : (d2 < sq_max);
}

template <int MODE>
__global__ void __launch_bounds__(kWarps * 32)
k_search(const float *__restrict__ xyzs, int64_t pn, const int32_t *__restrict__ block_begs,
         const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes, int64_t n_super,
         float sq_min, float sq_max, float cull_r, int use_min,
         int fixed_size, const int32_t *__restrict__ begs, int32_t *__restrict__ lens, int32_t *__restrict__ idxs,
         int32_t *__restrict__ cens) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t bb = 0, be = pn;
    if (block_begs) { bb = block_begs[blockIdx.y]; be = block_begs[blockIdx.y + 1]; }
    // max_block_pn (the grid's x extent) smaller than this block: the uncovered tail centres get empty rows instead of
    // uninitialised lengths (the contract is max_block_pn >= the largest block; this keeps a stale value harmless)
    if (MODE == MODE_COUNT && blockIdx.x == 0 && warp == 0)
        for (int64_t i = bb + (int64_t)gridDim.x * kCentres + lane; i < be; i += 32) lens[i] = 0;
    const int64_t c0 = bb + (int64_t)blockIdx.x * kCentres + warp * kCPW;
    if (c0 >= be) return;  // whole warp; warps are independent (no block-level barrier below)

    float cx[kCPW], cy[kCPW], cz[kCPW];
    int cnt[kCPW];
    int64_t wpos[kCPW];
    int last[kCPW];  // MODE_FIXED: last hit written
    const float inf = __int_as_float(0x7f800000);
    float blo[3] = {inf, inf, inf}, bhi[3] = {-inf, -inf, -inf};  // box of this warp's centres
#pragma unroll
    for (int c = 0; c < kCPW; ++c) {
        const int64_t ci = c0 + c;
        const bool v = ci < be;
        cx[c] = v ? __ldg(xyzs + ci * 3) : 0.f;
        cy[c] = v ? __ldg(xyzs + ci * 3 + 1) : 0.f;
        cz[c] = v ? __ldg(xyzs + ci * 3 + 2) : 0.f;
        if (v) {
            blo[0] = fminf(blo[0], cx[c]); bhi[0] = fmaxf(bhi[0], cx[c]);
            blo[1] = fminf(blo[1], cy[c]); bhi[1] = fmaxf(bhi[1], cy[c]);
            blo[2] = fminf(blo[2], cz[c]); bhi[2] = fmaxf(bhi[2], cz[c]);
        }
        cnt[c] = 0;
        wpos[c] = 0;
        last[c] = (int)ci;
        if (MODE == MODE_FILL && v) wpos[c] = begs[ci];
        if (MODE == MODE_FIXED) wpos[c] = ci * fixed_size;
    }
    const unsigned lt = (1u << lane) - 1u;
    const int64_t g_lo = bb >> 5, g_hi = (be + 31) >> 5;  // chunks touching this block
    const int64_t s_lo = g_lo >> 5, s_hi = (g_hi + 31) >> 5;  // super-chunks touching this block
    bool all_done = false;
    for (int64_t s0 = s_lo; s0 < s_hi && !all_done; s0 += 32) {
      // lane = super-chunk: conservative box-vs-box test.  (gap > r) is false for NaN => never culled.
      const int64_t sidx = s0 + lane;
      bool keep2 = false;
      if (sidx < s_hi) {
          keep2 = true;
#pragma unroll
          for (int k = 0; k < 3; ++k) {
              const float clo = __ldg(sboxes + k * n_super + sidx), chi = __ldg(sboxes + (3 + k) * n_super + sidx);
              if (__fsub_rn(clo, bhi[k]) > cull_r || __fsub_rn(blo[k], chi) > cull_r) keep2 = false;
          }
      }
      unsigned todo2 = __ballot_sync(0xffffffffu, keep2);
      while (todo2 && !all_done) {
        const int64_t g0 = (s0 + (__ffs(todo2) - 1)) * 32;
        todo2 &= todo2 - 1;
        // lane = chunk of the surviving super-chunk
        const int64_t g = g0 + lane;
        bool keep = false;
        if (g >= g_lo && g < g_hi) {
            keep = true;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float clo = __ldg(boxes + k * n_chunks + g), chi = __ldg(boxes + (3 + k) * n_chunks + g);
                if (__fsub_rn(clo, bhi[k]) > cull_r || __fsub_rn(blo[k], chi) > cull_r) keep = false;
            }
        }
        unsigned todo = __ballot_sync(0xffffffffu, keep);
        while (todo) {
            const int b = __ffs(todo) - 1;
            todo &= todo - 1;
            const int64_t j = (g0 + b) * 32 + lane;
            const bool jv = j >= bb && j < be;
            float tx = 0.f, ty = 0.f, tz = 0.f;
            if (jv) { tx = __ldg(xyzs + j * 3); ty = __ldg(xyzs + j * 3 + 1); tz = __ldg(xyzs + j * 3 + 2); }
#pragma unroll
            for (int c = 0; c < kCPW; ++c) {
                const bool h = jv && (c0 + c < be) && hit_test(tx, ty, tz, cx[c], cy[c], cz[c], sq_min, sq_max, use_min);
                const unsigned bal = __ballot_sync(0xffffffffu, h);
                if (MODE == MODE_COUNT) {
                    cnt[c] += __popc(bal);
                } else if (MODE == MODE_FILL) {
                    if (h) {
                        const int64_t w = wpos[c] + __popc(bal & lt);
                        idxs[w] = (int32_t)j;
                        cens[w] = (int32_t)(c0 + c);
                    }
                    wpos[c] += __popc(bal);
                } else {
                    const int slot = cnt[c] + __popc(bal & lt);
                    if (h && slot < fixed_size) {
                        idxs[wpos[c] + slot] = (int32_t)j;
                        cens[wpos[c] + slot] = (int32_t)(c0 + c);
                    }
                    const int room = fixed_size - cnt[c];  // last hit that made it into the row (warp-uniform)
                    if (bal && room > 0) {
                        unsigned b2 = bal;
                        const int nh = __popc(bal);
                        for (int k = 0; k < nh - room; ++k) b2 &= ~(1u << (31 - __clz(b2)));
                        last[c] = (int)((g0 + b) * 32 + (31 - __clz(b2)));
                    }
                    cnt[c] += __popc(bal);
                }
            }
            if (MODE == MODE_FIXED) {
                // early exit once every centre of the warp has its K hits (reference: loop condition
                // counts<fixed_size, SearchNeighborhoodFixed.cu:26)
                bool done = true;
#pragma unroll
                for (int c = 0; c < kCPW; ++c) done = done && (c0 + c >= be || cnt[c] >= fixed_size);
                if (done) { all_done = true; break; }
            }
        }
      }
    }
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

extern "C" size_t pcs_search_workspace_bytes(int64_t pn) {
    return boxes_bytes(pn > 0 ? pn : 1) + pcs_exclusive_scan_workspace_bytes(pn);
}

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
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_search<MODE_COUNT><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, s>>>(
        xyzs, pn, block_begs, boxes, n_chunks_of(pn), sboxes, n_super_of(pn), sq_min, sq_max, cull_radius(sq_max), use_min, 0, nullptr, lens,
        nullptr, nullptr);
    if (check_launch("search_count")) return 1;
    return pcs_exclusive_scan_i32(lens, pn, begs, total_dev, static_cast<char *>(workspace) + boxes_bytes(pn),
                                  workspace_bytes - boxes_bytes(pn), stream);
}

// `workspace` may be the one pcs_search_neighborhood_count was given (the chunk boxes are rebuilt
// here, so it need not have been preserved).
extern "C" int pcs_search_neighborhood_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                            int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                            const int32_t *begs, int32_t *idxs, int32_t *cens, void *workspace,
                                            size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_fill: bad pn");
    if (pn == 0) return 0;
    if (check_blocks(block_begs, nb, max_block_pn, pn)) return 1;
    cudaStream_t s = as_stream(stream);
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_search<MODE_FILL><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, s>>>(
        xyzs, pn, block_begs, boxes, n_chunks_of(pn), sboxes, n_super_of(pn), sq_min, sq_max, cull_radius(sq_max), use_min, 0, begs, nullptr,
        idxs, cens);
    return check_launch("search_fill");
}

extern "C" int pcs_search_neighborhood_fixed(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb,
                                             int64_t max_block_pn, float sq_min, float sq_max, int use_min,
                                             int64_t fixed_size, int32_t *idxs, int32_t *lens, int32_t *begs,
                                             int32_t *cens, void *workspace, size_t workspace_bytes,
                                             pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && fixed_size >= 0 && pn * fixed_size < ((int64_t)1 << 31), "search_fixed: bad size");
    if (pn == 0) return 0;
    if (check_blocks(block_begs, nb, max_block_pn, pn)) return 1;
    cudaStream_t s = as_stream(stream);
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_search<MODE_FIXED><<<search_grid(pn, nb, max_block_pn, block_begs != nullptr), kWarps * 32, 0, s>>>(
        xyzs, pn, block_begs, boxes, n_chunks_of(pn), sboxes, n_super_of(pn), sq_min, sq_max, cull_radius(sq_max), use_min, (int)fixed_size,
        nullptr, lens, idxs, cens);
    if (check_launch("search_fixed")) return 1;
    k_fixed_begs<<<(pn + 255) / 256, 256, 0, s>>>(begs, pn, (int)fixed_size);
    return check_launch("fixed_begs");
}

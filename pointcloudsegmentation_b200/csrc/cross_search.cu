// cross_search.cu — neighbour construction between TWO point sets (the libPointUtil calls the
// reference makes either side of the hot path; libPointUtil is un-vendored, semantics from the call sites):
//   * radius search of queries Q into sources S      (findNeighborRadiusCPU(xyzs, ds_idxs, r), s3dis_util.py:66)
//   * exact K nearest sources of every query         (findNeighborInAnotherCPU(sxyzs, qxyzs, K), interpolate.py:136)
// Same machinery as search.cu: sources visited in ascending index order in 32-point chunks (lane = candidate), two
// levels of chunk bounding boxes, a warp owns 8 consecutive queries and culls chunks against the queries' box
// inflated by the search radius.  Rows come out in ascending source index (radius search) or ascending
// (distance, index) (k-NN).  Predicate: d2 = fma(dz,dz, fma(dx,dx, dy*dy)), d = source - query, strict `<`, exactly as
// the self search (SearchNeighborhood.cu:13-19).
//
// k-NN = a radius per query that is known to hold >= K sources, then a top-K selection inside it:
//   pcs_knn_radius  — one round: every query that has no radius yet counts the sources with d2 < r2 (early exit at K);
//                     if it finds K it records r2.  The host issues rounds with doubling r, then one with r = inf
//                     (no host sync in between: satisfied warps return at once).
//   pcs_knn_select  — per query, the K smallest (d2, index) among the sources with d2 < its r2: lane k of the warp
//                     holds entry k of the sorted list, a hit is inserted with one ballot + shuffle-up.
#include "common.cuh"
#include "search_boxes.cuh"

using namespace pcs;

namespace {
constexpr int kWarps = 8;
constexpr int kCPW = 8;  // queries per warp
constexpr int kQ = kWarps * kCPW;

__device__ __forceinline__ float dist2(float tx, float ty, float tz, float cx, float cy, float cz) {
    const float dx = __fsub_rn(tx, cx), dy = __fsub_rn(ty, cy), dz = __fsub_rn(tz, cz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

struct Queries {
    float cx[kCPW], cy[kCPW], cz[kCPW];
    float blo[3], bhi[3];
};

__device__ __forceinline__ void load_queries(const float *__restrict__ q, int64_t c0, int64_t qn, Queries &Q) {
    const float inf = __int_as_float(0x7f800000);
#pragma unroll
    for (int k = 0; k < 3; ++k) { Q.blo[k] = inf; Q.bhi[k] = -inf; }
#pragma unroll
    for (int c = 0; c < kCPW; ++c) {
        const int64_t ci = c0 + c;
        const bool v = ci < qn;
        Q.cx[c] = v ? __ldg(q + ci * 3) : 0.f;
        Q.cy[c] = v ? __ldg(q + ci * 3 + 1) : 0.f;
        Q.cz[c] = v ? __ldg(q + ci * 3 + 2) : 0.f;
        if (v) {
            Q.blo[0] = fminf(Q.blo[0], Q.cx[c]); Q.bhi[0] = fmaxf(Q.bhi[0], Q.cx[c]);
            Q.blo[1] = fminf(Q.blo[1], Q.cy[c]); Q.bhi[1] = fmaxf(Q.bhi[1], Q.cy[c]);
            Q.blo[2] = fminf(Q.blo[2], Q.cz[c]); Q.bhi[2] = fmaxf(Q.bhi[2], Q.cz[c]);
        }
    }
}

// Visit every 32-source chunk whose box is within cull_r of the queries' box, in ascending order; f(first source
// index of the chunk) returns true to stop.  Whole-warp; cull_r = inf or NaN visits everything.
template <class F>
__device__ __forceinline__ void for_each_chunk(const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes,
                                               int64_t n_super, const Queries &Q, float cull_r, F f) {
    const int lane = threadIdx.x & 31;
    for (int64_t s0 = 0; s0 < n_super; s0 += 32) {
        const int64_t sidx = s0 + lane;
        bool keep2 = false;
        if (sidx < n_super) {
            keep2 = true;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const float clo = __ldg(sboxes + k * n_super + sidx), chi = __ldg(sboxes + (3 + k) * n_super + sidx);
                if (__fsub_rn(clo, Q.bhi[k]) > cull_r || __fsub_rn(Q.blo[k], chi) > cull_r) keep2 = false;
            }
        }
        unsigned todo2 = __ballot_sync(0xffffffffu, keep2);
        while (todo2) {
            const int64_t g0 = (s0 + (__ffs(todo2) - 1)) * 32;
            todo2 &= todo2 - 1;
            const int64_t g = g0 + lane;
            bool keep = false;
            if (g < n_chunks) {
                keep = true;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float clo = __ldg(boxes + k * n_chunks + g), chi = __ldg(boxes + (3 + k) * n_chunks + g);
                    if (__fsub_rn(clo, Q.bhi[k]) > cull_r || __fsub_rn(Q.blo[k], chi) > cull_r) keep = false;
                }
            }
            unsigned todo = __ballot_sync(0xffffffffu, keep);
            while (todo) {
                const int b = __ffs(todo) - 1;
                todo &= todo - 1;
                if (f((g0 + b) * 32)) return;
            }
        }
    }
}

// ---- radius search Q -> S: FILL = false counts (lens), FILL = true writes rows at begs -------------------------
template <bool FILL>
__global__ void __launch_bounds__(kWarps * 32)
k_cross_search(const float *__restrict__ sxyzs, int64_t sn, const float *__restrict__ qxyzs, int64_t qn,
               const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes, int64_t n_super,
               float sq_min, float sq_max, float cull_r, int use_min, const int32_t *__restrict__ begs,
               int32_t *__restrict__ lens, int32_t *__restrict__ idxs, int32_t *__restrict__ cens) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c0 = (int64_t)blockIdx.x * kQ + warp * kCPW;
    if (c0 >= qn) return;
    Queries Q;
    load_queries(qxyzs, c0, qn, Q);
    int cnt[kCPW];
    int64_t wpos[kCPW];
#pragma unroll
    for (int c = 0; c < kCPW; ++c) { cnt[c] = 0; wpos[c] = (FILL && c0 + c < qn) ? begs[c0 + c] : 0; }
    const unsigned lt = (1u << lane) - 1u;
    for_each_chunk(boxes, n_chunks, sboxes, n_super, Q, cull_r, [&](int64_t j0) {
        const int64_t j = j0 + lane;
        const bool jv = j < sn;
        float tx = 0.f, ty = 0.f, tz = 0.f;
        if (jv) { tx = __ldg(sxyzs + j * 3); ty = __ldg(sxyzs + j * 3 + 1); tz = __ldg(sxyzs + j * 3 + 2); }
#pragma unroll
        for (int c = 0; c < kCPW; ++c) {
            const float d2 = dist2(tx, ty, tz, Q.cx[c], Q.cy[c], Q.cz[c]);
            const bool h = jv && (c0 + c < qn) && (use_min ? (d2 < sq_max && d2 > sq_min) : (d2 < sq_max));
            const unsigned bal = __ballot_sync(0xffffffffu, h);
            if (FILL) {
                if (h) {
                    const int64_t w = wpos[c] + __popc(bal & lt);
                    idxs[w] = (int32_t)j;
                    cens[w] = (int32_t)(c0 + c);
                }
                wpos[c] += __popc(bal);
            } else {
                cnt[c] += __popc(bal);
            }
        }
        return false;
    });
    if (!FILL) {
#pragma unroll
        for (int c = 0; c < kCPW; ++c)
            if (lane == c && c0 + c < qn) lens[c0 + c] = cnt[c];
    }
}

__global__ void k_fill_f32(float *__restrict__ p, int64_t n, float v) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- k-NN, step 1: one radius round ---------------------------------------------------------------------------
// rad2[q] < 0 = no radius yet.  A query that finds >= K sources with d2 < r2 records rad2[q] = r2.
__global__ void __launch_bounds__(kWarps * 32)
k_knn_radius(const float *__restrict__ sxyzs, int64_t sn, const float *__restrict__ qxyzs, int64_t qn,
             const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes, int64_t n_super,
             float r2, float cull_r, int K, float *__restrict__ rad2) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c0 = (int64_t)blockIdx.x * kQ + warp * kCPW;
    if (c0 >= qn) return;
    bool need[kCPW];
    bool any = false;
#pragma unroll
    for (int c = 0; c < kCPW; ++c) { need[c] = (c0 + c < qn) && (__ldg(rad2 + c0 + c) < 0.f); any = any || need[c]; }
    if (!any) return;  // warp-uniform: every lane read the same eight values
    Queries Q;
    load_queries(qxyzs, c0, qn, Q);
    int cnt[kCPW];
#pragma unroll
    for (int c = 0; c < kCPW; ++c) cnt[c] = 0;
    for_each_chunk(boxes, n_chunks, sboxes, n_super, Q, cull_r, [&](int64_t j0) {
        const int64_t j = j0 + lane;
        const bool jv = j < sn;
        float tx = 0.f, ty = 0.f, tz = 0.f;
        if (jv) { tx = __ldg(sxyzs + j * 3); ty = __ldg(sxyzs + j * 3 + 1); tz = __ldg(sxyzs + j * 3 + 2); }
        bool done = true;
#pragma unroll
        for (int c = 0; c < kCPW; ++c) {
            const bool h = jv && need[c] && dist2(tx, ty, tz, Q.cx[c], Q.cy[c], Q.cz[c]) < r2;
            cnt[c] += __popc(__ballot_sync(0xffffffffu, h));
            done = done && (!need[c] || cnt[c] >= K);
        }
        return done;
    });
#pragma unroll
    for (int c = 0; c < kCPW; ++c)
        if (lane == c && need[c] && cnt[c] >= K) rad2[c0 + c] = r2;
}

// ---- k-NN, step 2: top-K selection inside each query's radius ----------------------------------------------------
// Lane k (< K <= 32) holds entry k of every query's list, sorted ascending by (d2, source index).
__global__ void __launch_bounds__(kWarps * 32)
k_knn_select(const float *__restrict__ sxyzs, int64_t sn, const float *__restrict__ qxyzs, int64_t qn,
             const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes, int64_t n_super,
             const float *__restrict__ rad2, int K, int32_t *__restrict__ idxs, float *__restrict__ dists) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t c0 = (int64_t)blockIdx.x * kQ + warp * kCPW;
    if (c0 >= qn) return;
    Queries Q;
    load_queries(qxyzs, c0, qn, Q);
    const float inf = __int_as_float(0x7f800000);
    float r2[kCPW], ld[kCPW], wd[kCPW];  // radius bound; this lane's entry; the list's current worst (entry K-1)
    int lj[kCPW], wj[kCPW];
    float rmax = 0.f;
#pragma unroll
    for (int c = 0; c < kCPW; ++c) {
        const float r = (c0 + c < qn) ? __ldg(rad2 + c0 + c) : 0.f;
        r2[c] = r < 0.f ? inf : r;  // no radius recorded (fewer than K sources in reach of any round): everything
        rmax = fmaxf(rmax, r2[c]);
        ld[c] = wd[c] = inf;
        lj[c] = wj[c] = 0x7fffffff;
    }
    const float cull_r = sqrtf(rmax) * 1.001f;
    const unsigned kmask = K >= 32 ? 0xffffffffu : ((1u << K) - 1u);
    for_each_chunk(boxes, n_chunks, sboxes, n_super, Q, cull_r, [&](int64_t j0) {
        const int64_t j = j0 + lane;
        const bool jv = j < sn;
        float tx = 0.f, ty = 0.f, tz = 0.f;
        if (jv) { tx = __ldg(sxyzs + j * 3); ty = __ldg(sxyzs + j * 3 + 1); tz = __ldg(sxyzs + j * 3 + 2); }
#pragma unroll
        for (int c = 0; c < kCPW; ++c) {
            const float d2 = dist2(tx, ty, tz, Q.cx[c], Q.cy[c], Q.cz[c]);
            // candidates arrive in ascending index, so (d2, j) beats the worst entry iff d2 < its distance
            // (an equal distance has a larger index than everything already in the list)
            bool h = jv && (c0 + c < qn) && d2 < r2[c] && d2 < wd[c];
            unsigned bal = __ballot_sync(0xffffffffu, h);
            while (bal) {
                const int b = __ffs(bal) - 1;
                bal &= bal - 1;
                const float nd = __shfl_sync(0xffffffffu, d2, b);
                const int nj = (int)(j0 + b);
                if (!(nd < wd[c])) continue;  // an earlier insertion of this ballot tightened the bound (warp-uniform)
                // entries strictly before the new one form a prefix of the sorted list
                const bool before = (ld[c] < nd) || (ld[c] == nd && lj[c] < nj);
                const int pos = __popc(__ballot_sync(0xffffffffu, before) & kmask);
                const float ud = __shfl_up_sync(0xffffffffu, ld[c], 1);
                const int uj = __shfl_up_sync(0xffffffffu, lj[c], 1);
                if (lane == pos) { ld[c] = nd; lj[c] = nj; }
                else if (lane > pos) { ld[c] = ud; lj[c] = uj; }
                wd[c] = __shfl_sync(0xffffffffu, ld[c], K - 1);
                wj[c] = __shfl_sync(0xffffffffu, lj[c], K - 1);
            }
        }
        return false;
    });
#pragma unroll
    for (int c = 0; c < kCPW; ++c) {
        if (c0 + c < qn && lane < K) {
            idxs[(c0 + c) * K + lane] = lj[c] == 0x7fffffff ? -1 : lj[c];
            if (dists) dists[(c0 + c) * K + lane] = ld[c];
        }
    }
}
}  // namespace

extern "C" size_t pcs_cross_search_workspace_bytes(int64_t sn, int64_t qn) {
    return boxes_bytes(sn > 0 ? sn : 1) + pcs_exclusive_scan_workspace_bytes(qn);
}

extern "C" int pcs_cross_search_count(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, float sq_min,
                                      float sq_max, int use_min, int32_t *lens, int32_t *begs, int32_t *total_dev,
                                      void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(sn >= 0 && qn >= 0 && sn < ((int64_t)1 << 31) && qn < ((int64_t)1 << 31), "cross_search_count: bad size");
    cudaStream_t s = as_stream(stream);
    if (qn == 0) { PCS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), s)); return 0; }
    if (sn == 0) {
        PCS_CUDA(cudaMemsetAsync(lens, 0, qn * sizeof(int32_t), s));
        PCS_CUDA(cudaMemsetAsync(begs, 0, qn * sizeof(int32_t), s));
        PCS_CUDA(cudaMemsetAsync(total_dev, 0, sizeof(int32_t), s));
        return 0;
    }
    float *boxes, *sboxes;
    if (make_boxes(sxyzs, sn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_cross_search<false><<<(unsigned)((qn + kQ - 1) / kQ), kWarps * 32, 0, s>>>(
        sxyzs, sn, qxyzs, qn, boxes, n_chunks_of(sn), sboxes, n_super_of(sn), sq_min, sq_max, cull_radius(sq_max), use_min,
        nullptr, lens, nullptr, nullptr);
    if (check_launch("cross_search_count")) return 1;
    return pcs_exclusive_scan_i32(lens, qn, begs, total_dev, static_cast<char *>(workspace) + boxes_bytes(sn),
                                  workspace_bytes - boxes_bytes(sn), stream);
}

extern "C" int pcs_cross_search_fill(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, float sq_min,
                                     float sq_max, int use_min, const int32_t *begs, int32_t *idxs, int32_t *cens,
                                     void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(sn >= 0 && qn >= 0 && sn < ((int64_t)1 << 31) && qn < ((int64_t)1 << 31), "cross_search_fill: bad size");
    if (qn == 0 || sn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    float *boxes, *sboxes;
    if (make_boxes(sxyzs, sn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_cross_search<true><<<(unsigned)((qn + kQ - 1) / kQ), kWarps * 32, 0, s>>>(
        sxyzs, sn, qxyzs, qn, boxes, n_chunks_of(sn), sboxes, n_super_of(sn), sq_min, sq_max, cull_radius(sq_max), use_min,
        begs, nullptr, idxs, cens);
    return check_launch("cross_search_fill");
}

extern "C" int pcs_knn_cross(const float *sxyzs, int64_t sn, const float *qxyzs, int64_t qn, int64_t K, float r0,
                             int64_t rounds, int32_t *idxs, float *dists /* nullable */, float *rad2 /* [qn] scratch */,
                             void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(sn >= 0 && qn >= 0 && sn < ((int64_t)1 << 31) && qn * K < ((int64_t)1 << 31), "knn_cross: bad size");
    PCS_REQUIRE(K >= 1 && K <= 32, "knn_cross: K must be in [1,32]");
    PCS_REQUIRE(sn >= K, "knn_cross: fewer sources (%lld) than K (%lld)", (long long)sn, (long long)K);
    PCS_REQUIRE(r0 > 0.f && rounds >= 0 && rounds <= 32, "knn_cross: bad r0 / rounds");
    if (qn == 0) return 0;
    cudaStream_t s = as_stream(stream);
    float *boxes, *sboxes;
    if (make_boxes(sxyzs, sn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    k_fill_f32<<<(unsigned)((qn + 255) / 256), 256, 0, s>>>(rad2, qn, -1.f);  // < 0 = no radius yet
    if (check_launch("knn_fill")) return 1;
    const unsigned grid = (unsigned)((qn + kQ - 1) / kQ);
    float r = r0;
    for (int64_t it = 0; it <= rounds; ++it) {
        const bool last = it == rounds;
        const float r2 = last ? INFINITY : r * r;
        k_knn_radius<<<grid, kWarps * 32, 0, s>>>(sxyzs, sn, qxyzs, qn, boxes, n_chunks_of(sn), sboxes, n_super_of(sn), r2,
                                                 last ? r2 : cull_radius(r2), (int)K, rad2);
        if (check_launch("knn_radius")) return 1;
        r *= 2.f;
    }
    k_knn_select<<<grid, kWarps * 32, 0, s>>>(sxyzs, sn, qxyzs, qn, boxes, n_chunks_of(sn), sboxes, n_super_of(sn), rad2, (int)K,
                                             idxs, dists);
    return check_launch("knn_select");
}

// search_multi.cu — several radius / annulus graphs over the SAME points from ONE traversal (a18/a19).
// The towers build up to four graphs per level (pointnet_13_dilated_embed level 0: ball 0.15, annuli 0.15-0.2 and
// 0.1-0.15, ball 0.1; model_pointnet.py:935-1000).  search.cu runs a full culled pass per graph, twice (count, fill);
// here a warp walks the chunk boxes once with the cull radius of the widest graph, computes every candidate's squared
// distance once and evaluates the NG predicates on it — the box tests, the coordinate loads and the distance (2/3 of
// the instructions of a pass) are shared.  Every graph's rows are exactly what search.cu writes for it (same
// predicate, same ascending order): tested bit-identical.  Host side: ONE count launch, NG scans, ONE sizing read for
// all graphs of the level (instead of one per graph), ONE fill launch.
// Round 2, one traversal instead of two: the count pass also records every hit in a padded list (pad[g][centre][maxk],
// the same ascending order); "fill" is then a row copy pad -> CSR, and only the warps owning a centre with more than maxk
// neighbours traverse again (pcs_search_multi_count_pad / pcs_search_multi_fill_pad).
#include "common.cuh"
#include "search_boxes.cuh"

using namespace pcs;

namespace {
constexpr int kWarps = 8;
constexpr int kCPW = 8;
constexpr int kCentres = kWarps * kCPW;
constexpr int kMaxG = 4;

struct MultiParams {
    float sq_min[kMaxG], sq_max[kMaxG];
    int use_min[kMaxG];
    int32_t *lens[kMaxG];        // count: out
    const int32_t *begs[kMaxG];  // fill: in
    int32_t *idxs[kMaxG], *cens[kMaxG];
    int32_t *pad[kMaxG];  // count: hit lists [pn][maxk] (nullable); fill with maxk > 0: only centres with lens > maxk are written
    int maxk;
};

template <bool FILL, int NG>
__global__ void __launch_bounds__(kWarps * 32)
k_search_multi(const float *__restrict__ xyzs, int64_t pn, const int32_t *__restrict__ block_begs,
               const float *__restrict__ boxes, int64_t n_chunks, const float *__restrict__ sboxes, int64_t n_super,
               float cull_r, MultiParams P) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int64_t bb = 0, be = pn;
    if (block_begs) { bb = block_begs[blockIdx.y]; be = block_begs[blockIdx.y + 1]; }
    // max_block_pn smaller than this block: uncovered tail centres get empty rows (see search.cu)
    if (!FILL && blockIdx.x == 0 && warp == 0)
        for (int64_t i = bb + (int64_t)gridDim.x * kCentres + lane; i < be; i += 32)
#pragma unroll
            for (int g = 0; g < NG; ++g) P.lens[g][i] = 0;
    const int64_t c0 = bb + (int64_t)blockIdx.x * kCentres + warp * kCPW;
    if (c0 >= be) return;
    const int maxk = P.maxk;
    bool over[kCPW][NG];  // fill after a padded count: this (centre, graph) row did not fit its padded list
    if (FILL && maxk > 0) {
        bool any = false;
#pragma unroll
        for (int c = 0; c < kCPW; ++c)
#pragma unroll
            for (int g = 0; g < NG; ++g) {
                over[c][g] = c0 + c < be && __ldg(P.lens[g] + c0 + c) > maxk;
                any |= over[c][g];
            }
        if (!any) return;  // (warp-uniform: every lane read the same counts)
    }
    float cx[kCPW], cy[kCPW], cz[kCPW];
    int acc[kCPW][NG];  // count: hits so far; fill: next write position (edge counts fit int32: en < 2^31)
    const float inf = __int_as_float(0x7f800000);
    float blo[3] = {inf, inf, inf}, bhi[3] = {-inf, -inf, -inf};
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
#pragma unroll
        for (int g = 0; g < NG; ++g) acc[c][g] = (FILL && v) ? P.begs[g][ci] : 0;
    }
    const unsigned lt = (1u << lane) - 1u;
    const int64_t g_lo = bb >> 5, g_hi = (be + 31) >> 5;
    const int64_t s_lo = g_lo >> 5, s_hi = (g_hi + 31) >> 5;
    for (int64_t s0 = s_lo; s0 < s_hi; s0 += 32) {
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
        while (todo2) {
            const int64_t g0 = (s0 + (__ffs(todo2) - 1)) * 32;
            todo2 &= todo2 - 1;
            const int64_t gc = g0 + lane;
            bool keep = false;
            if (gc >= g_lo && gc < g_hi) {
                keep = true;
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const float clo = __ldg(boxes + k * n_chunks + gc), chi = __ldg(boxes + (3 + k) * n_chunks + gc);
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
                    const float dx = __fsub_rn(tx, cx[c]), dy = __fsub_rn(ty, cy[c]), dz = __fsub_rn(tz, cz[c]);
                    const float d2 = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
                    const bool cv = jv && (c0 + c < be);
#pragma unroll
                    for (int g = 0; g < NG; ++g) {
                        const bool h = cv && d2 < P.sq_max[g] && (!P.use_min[g] || d2 > P.sq_min[g]);
                        const unsigned bal = __ballot_sync(0xffffffffu, h);
                        if (FILL && h && (maxk <= 0 || over[c][g])) {
                            const int w = acc[c][g] + __popc(bal & lt);
                            P.idxs[g][w] = (int32_t)j;
                            P.cens[g][w] = (int32_t)(c0 + c);
                        }
                        if (!FILL && maxk > 0 && h) {
                            const int w = acc[c][g] + __popc(bal & lt);
                            if (w < maxk) P.pad[g][(c0 + c) * maxk + w] = (int32_t)j;
                        }
                        acc[c][g] += __popc(bal);
                    }
                }
            }
        }
    }
    if (!FILL) {
#pragma unroll
        for (int c = 0; c < kCPW; ++c)
#pragma unroll
            for (int g = 0; g < NG; ++g)
                if (lane == c && c0 + c < be) P.lens[g][c0 + c] = acc[c][g];
    }
}

template <bool FILL>
int launch_multi(int ng, dim3 grid, cudaStream_t s, const float *xyzs, int64_t pn, const int32_t *block_begs, const float *boxes,
                 const float *sboxes, float cull_r, const MultiParams &P) {
    const int64_t nc = n_chunks_of(pn), ns = n_super_of(pn);
    switch (ng) {
        case 1: k_search_multi<FILL, 1><<<grid, kWarps * 32, 0, s>>>(xyzs, pn, block_begs, boxes, nc, sboxes, ns, cull_r, P); break;
        case 2: k_search_multi<FILL, 2><<<grid, kWarps * 32, 0, s>>>(xyzs, pn, block_begs, boxes, nc, sboxes, ns, cull_r, P); break;
        case 3: k_search_multi<FILL, 3><<<grid, kWarps * 32, 0, s>>>(xyzs, pn, block_begs, boxes, nc, sboxes, ns, cull_r, P); break;
        default: k_search_multi<FILL, 4><<<grid, kWarps * 32, 0, s>>>(xyzs, pn, block_begs, boxes, nc, sboxes, ns, cull_r, P); break;
    }
    return check_launch(FILL ? "search_multi_fill" : "search_multi_count");
}

int fill_params(MultiParams &P, int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min, float *cull_r) {
    PCS_REQUIRE(ng >= 1 && ng <= kMaxG, "search_multi: between 1 and %d graphs per call", kMaxG);
    float rmax = 0.f;
    for (int g = 0; g < kMaxG; ++g) {
        const int k = g < ng ? g : 0;
        P.sq_min[g] = sq_min[k]; P.sq_max[g] = sq_max[k]; P.use_min[g] = use_min[k];
        P.lens[g] = nullptr; P.begs[g] = nullptr; P.idxs[g] = nullptr; P.cens[g] = nullptr; P.pad[g] = nullptr;
        if (g < ng) rmax = (sq_max[g] > rmax || sq_max[g] != sq_max[g]) ? sq_max[g] : rmax;  // NaN -> no culling
    }
    P.maxk = 0;
    *cull_r = cull_radius(rmax);
    return 0;
}

// rows that fit their padded list: pad[centre][0 .. len) -> idxs[beg ..], cens[beg ..] = centre.  A warp takes 32 consecutive
// centres: one coalesced read of their counts and offsets, then 32 independent row copies (a warp per centre was one
// dependent chain of three memory latencies per centre: 0.7 ms per call).
__global__ void __launch_bounds__(256)
k_search_compact(const int32_t *__restrict__ lens, const int32_t *__restrict__ begs, const int32_t *__restrict__ pad, int64_t pn,
                 int maxk, int32_t *__restrict__ idxs, int32_t *__restrict__ cens) {
    const int lane = threadIdx.x & 31;
    const int64_t W = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t grp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); grp * 32 < pn; grp += W) {
        const int64_t i = grp * 32 + lane;
        int len = i < pn ? __ldg(lens + i) : 0;
        const int beg = i < pn ? __ldg(begs + i) : 0;
        if (len > maxk) len = 0;  // written by the second traversal
#pragma unroll 4
        for (int c = 0; c < 32; ++c) {
            const int lc = __shfl_sync(0xffffffffu, len, c), bc = __shfl_sync(0xffffffffu, beg, c);
            const int64_t ic = grp * 32 + c;
            for (int k = lane; k < lc; k += 32) {
                idxs[bc + k] = __ldg(pad + ic * maxk + k);
                cens[bc + k] = (int32_t)ic;
            }
        }
    }
}

dim3 multi_grid(int64_t pn, int64_t nb, int64_t max_block_pn, bool batched) {
    const int64_t per = batched ? max_block_pn : pn;
    return dim3((unsigned)((per + kCentres - 1) / kCentres), (unsigned)(batched ? nb : 1), 1);
}
}  // namespace

extern "C" int pcs_search_multi_count(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                                      int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min,
                                      int32_t *lens, int32_t *begs, int32_t *totals_dev, void *workspace,
                                      size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_multi_count: bad pn");
    MultiParams P;
    float cull_r;
    if (fill_params(P, ng, sq_min, sq_max, use_min, &cull_r)) return 1;
    cudaStream_t s = as_stream(stream);
    if (pn == 0) { PCS_CUDA(cudaMemsetAsync(totals_dev, 0, ng * sizeof(int32_t), s)); return 0; }
    if (block_begs) {
        PCS_REQUIRE(nb >= 1 && nb <= 65535, "search_multi: nb must be in [1,65535]");
        PCS_REQUIRE(max_block_pn >= 1 && max_block_pn <= pn, "search_multi: bad max_block_pn");
    }
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    for (int g = 0; g < ng; ++g) P.lens[g] = lens + g * pn;
    if (launch_multi<false>((int)ng, multi_grid(pn, nb, max_block_pn, block_begs != nullptr), s, xyzs, pn, block_begs, boxes, sboxes,
                            cull_r, P)) return 1;
    for (int g = 0; g < ng; ++g)
        if (pcs_exclusive_scan_i32(lens + g * pn, pn, begs + g * pn, totals_dev + g, static_cast<char *>(workspace) + boxes_bytes(pn),
                                   workspace_bytes - boxes_bytes(pn), stream)) return 1;
    return 0;
}

extern "C" int pcs_search_multi_fill(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                                     int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min,
                                     const int32_t *begs, int32_t *const *idxs, int32_t *const *cens, void *workspace,
                                     size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_multi_fill: bad pn");
    MultiParams P;
    float cull_r;
    if (fill_params(P, ng, sq_min, sq_max, use_min, &cull_r)) return 1;
    if (pn == 0) return 0;
    if (block_begs) {
        PCS_REQUIRE(nb >= 1 && nb <= 65535, "search_multi: nb must be in [1,65535]");
        PCS_REQUIRE(max_block_pn >= 1 && max_block_pn <= pn, "search_multi: bad max_block_pn");
    }
    cudaStream_t s = as_stream(stream);
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    for (int g = 0; g < ng; ++g) { P.begs[g] = begs + g * pn; P.idxs[g] = idxs[g]; P.cens[g] = cens[g]; }
    // graphs beyond ng alias graph 0's predicate but are never instantiated (NG = ng)
    return launch_multi<true>((int)ng, multi_grid(pn, nb, max_block_pn, block_begs != nullptr), s, xyzs, pn, block_begs, boxes, sboxes,
                              cull_r, P);
}

// ---- one traversal: count + padded hit lists, then a row copy ----------------------------------------------------
extern "C" int pcs_search_multi_count_pad(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                                          int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min,
                                          int32_t *lens, int32_t *begs, int32_t *totals_dev, int32_t *pad, int64_t maxk,
                                          void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_multi_count_pad: bad pn");
    PCS_REQUIRE(maxk >= 1 && maxk <= 4096 && pad != nullptr, "search_multi_count_pad: maxk must be in [1,4096] and pad non-null");
    MultiParams P;
    float cull_r;
    if (fill_params(P, ng, sq_min, sq_max, use_min, &cull_r)) return 1;
    cudaStream_t s = as_stream(stream);
    if (pn == 0) { PCS_CUDA(cudaMemsetAsync(totals_dev, 0, ng * sizeof(int32_t), s)); return 0; }
    if (block_begs) {
        PCS_REQUIRE(nb >= 1 && nb <= 65535, "search_multi: nb must be in [1,65535]");
        PCS_REQUIRE(max_block_pn >= 1 && max_block_pn <= pn, "search_multi: bad max_block_pn");
    }
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    for (int g = 0; g < ng; ++g) { P.lens[g] = lens + g * pn; P.pad[g] = pad + g * pn * maxk; }
    P.maxk = (int)maxk;
    if (launch_multi<false>((int)ng, multi_grid(pn, nb, max_block_pn, block_begs != nullptr), s, xyzs, pn, block_begs, boxes, sboxes,
                            cull_r, P)) return 1;
    for (int g = 0; g < ng; ++g)
        if (pcs_exclusive_scan_i32(lens + g * pn, pn, begs + g * pn, totals_dev + g, static_cast<char *>(workspace) + boxes_bytes(pn),
                                   workspace_bytes - boxes_bytes(pn), stream)) return 1;
    return 0;
}

extern "C" int pcs_search_multi_fill_pad(const float *xyzs, int64_t pn, const int32_t *block_begs, int64_t nb, int64_t max_block_pn,
                                         int64_t ng, const float *sq_min, const float *sq_max, const int32_t *use_min,
                                         const int32_t *lens, const int32_t *begs, const int32_t *pad, int64_t maxk,
                                         int32_t *const *idxs, int32_t *const *cens, void *workspace, size_t workspace_bytes,
                                         pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "search_multi_fill_pad: bad pn");
    PCS_REQUIRE(maxk >= 1 && maxk <= 4096 && pad != nullptr && lens != nullptr, "search_multi_fill_pad: bad padded lists");
    MultiParams P;
    float cull_r;
    if (fill_params(P, ng, sq_min, sq_max, use_min, &cull_r)) return 1;
    if (pn == 0) return 0;
    if (block_begs) {
        PCS_REQUIRE(nb >= 1 && nb <= 65535, "search_multi: nb must be in [1,65535]");
        PCS_REQUIRE(max_block_pn >= 1 && max_block_pn <= pn, "search_multi: bad max_block_pn");
    }
    cudaStream_t s = as_stream(stream);
    const int64_t want = (pn + 255) / 256;  // 32 centres per warp, 8 warps per CTA
    const unsigned cgrid = (unsigned)(want < 148 * 16 ? want : 148 * 16);
    for (int g = 0; g < ng; ++g) {
        k_search_compact<<<cgrid, 256, 0, s>>>(lens + g * pn, begs + g * pn, pad + g * pn * maxk, pn, (int)maxk, idxs[g], cens[g]);
        if (int e = check_launch("search_multi_compact")) return e;
    }
    // the rows that did not fit: a second traversal by the warps that own one (the others return after reading their counts)
    float *boxes, *sboxes;
    if (make_boxes(xyzs, pn, workspace, workspace_bytes, s, &boxes, &sboxes)) return 1;
    for (int g = 0; g < ng; ++g) {
        P.lens[g] = const_cast<int32_t *>(lens) + g * pn; P.begs[g] = begs + g * pn; P.idxs[g] = idxs[g]; P.cens[g] = cens[g];
    }
    P.maxk = (int)maxk;
    return launch_multi<true>((int)ng, multi_grid(pn, nb, max_block_pn, block_begs != nullptr), s, xyzs, pn, block_begs, boxes, sboxes,
                              cull_r, P);
}

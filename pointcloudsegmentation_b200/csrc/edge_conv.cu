// edge_conv.cu — the edge-tensor assembly of the split edge convolution (SURVEY.md §8f-1 extension;
// no reference CUDA counterpart: upstream does this with concat + fully_connected on [csum, 2C+3]
// tensors, model_pointnet.py:10-24 over graph_concat_scatter, tf_ops/graph_layer_new.py).
//
//   pcs_edge_combine_forward   u[e,:] = pa[cen[e],:] + pb[nbr[e],:] + Wx * sxyz[e]          [csum, H]
//   pcs_edge_xyz_wgrad         partial[b,h,c] = sum over CTA b's edges of du[e,h] * sxyz[e,c]
//
// pa / pb are the POINT-level products feats*Wc^T + bias and feats*Wn^T; the kernel replaces
// graph_unpool + neighbor_scatter + add + a K=3 GEMM (four passes over the [csum, H] tensor) by one
// flat pass that writes u once with STG.128.  Both kernels are HBM-bound on the [csum, H] tensor.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;
constexpr int kU = 4;       // independent 16-byte pieces per thread
constexpr int kMaxH4 = 64;  // H <= 256

template <typename IDX>
__global__ void __launch_bounds__(kBlock)
k_edge_combine(const float4 *__restrict__ pa, const float4 *__restrict__ pb, const float *__restrict__ sxyz,
               const float *__restrict__ wx, const int32_t *__restrict__ cen, const int32_t *__restrict__ nbr,
               IDX n_pieces, IDX H4, IDX ld4, float4 *__restrict__ u) {
    __shared__ float4 swx[3][kMaxH4];  // swx[c][q] = Wx[4q..4q+3][c]
    for (int t = threadIdx.x; t < 3 * (int)H4; t += kBlock) {
        const int c = t / (int)H4, q = t - c * (int)H4;
        swx[c][q] = make_float4(__ldg(wx + (4 * q + 0) * 3 + c), __ldg(wx + (4 * q + 1) * 3 + c),
                                __ldg(wx + (4 * q + 2) * 3 + c), __ldg(wx + (4 * q + 3) * 3 + c));
    }
    __syncthreads();
    const IDX base = (IDX)blockIdx.x * (kBlock * kU) + threadIdx.x;
    IDX row[kU], q[kU];
    int32_t ce[kU], nb[kU];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        row[k] = i / H4;
        q[k] = i - row[k] * H4;
        ce[k] = 0; nb[k] = 0;
        if (i < n_pieces) { ce[k] = __ldg(cen + row[k]); nb[k] = __ldg(nbr + row[k]); }
    }
    float4 a[kU], b[kU];
    float x[kU][3];
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            a[k] = __ldg(pa + (int64_t)ce[k] * ld4 + q[k]);
            b[k] = __ldg(pb + (int64_t)nb[k] * ld4 + q[k]);
            const float *p = sxyz + (int64_t)row[k] * 3;
            x[k][0] = __ldg(p); x[k][1] = __ldg(p + 1); x[k][2] = __ldg(p + 2);
        }
    }
#pragma unroll
    for (int k = 0; k < kU; ++k) {
        const IDX i = base + (IDX)k * kBlock;
        if (i < n_pieces) {
            const float4 w0 = swx[0][q[k]], w1 = swx[1][q[k]], w2 = swx[2][q[k]];
            float4 o;
            // (pa + pb) + ((w0*x + w1*y) + w2*z): a fixed order, so forward and the backward
            // recomputation see the same bits.
            o.x = (a[k].x + b[k].x) + fmaf(w2.x, x[k][2], fmaf(w1.x, x[k][1], w0.x * x[k][0]));
            o.y = (a[k].y + b[k].y) + fmaf(w2.y, x[k][2], fmaf(w1.y, x[k][1], w0.y * x[k][0]));
            o.z = (a[k].z + b[k].z) + fmaf(w2.z, x[k][2], fmaf(w1.z, x[k][1], w0.z * x[k][0]));
            o.w = (a[k].w + b[k].w) + fmaf(w2.w, x[k][2], fmaf(w1.w, x[k][1], w0.w * x[k][0]));
            u[i] = o;  // read again by the edge MLP right after: default cache policy
        }
    }
}

// Centre-walking form of k_edge_combine: a group of H/4 lanes owns one centre, keeps pa[centre] and its 4x3
// slice of Wx in registers and walks the centre's edges (ragged head first, then batches of 4): one index ->
// gather chain per 4 rows and no per-piece centre gather (the same trade as k_neighbor_scatter vs the flat
// kernel: 84-98 % vs 43-62 % of HBM peak for use_diff rows, tools/ab_rows.py).  Same arithmetic order as
// k_edge_combine, so both produce identical bits.
__global__ void __launch_bounds__(kBlock)  // (60 registers: 4 CTAs per SM; forcing 5 / 6 spills and costs 4 % / 25 %)
k_edge_combine_walk(const float *__restrict__ pa, const float *__restrict__ pb, const float *__restrict__ sxyz,
                    const float *__restrict__ wx, const int32_t *__restrict__ nidxs, const int32_t *__restrict__ lens,
                    const int32_t *__restrict__ bgs, int64_t pn, int64_t H, int64_t ld, int gshift, int CH,
                    float *__restrict__ u) {
    int64_t i; int col;
    if (!row_item(pn, gshift, CH, 4, H, i, col)) return;
    const int64_t bg = bgs[i];
    const int len = lens[i];
    const float4 a = __ldg(reinterpret_cast<const float4 *>(pa + i * ld + col));
    float w[4][3];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        w[k][0] = __ldg(wx + (col + k) * 3); w[k][1] = __ldg(wx + (col + k) * 3 + 1); w[k][2] = __ldg(wx + (col + k) * 3 + 2);
    }
    const int32_t *np = nidxs + bg;
    const float *sp = sxyz + bg * 3;
    float *op = u + bg * H + col;
    auto emit = [&](float *o, const float4 &b, float x, float y, float z) {
        float4 r;
        r.x = (a.x + b.x) + fmaf(w[0][2], z, fmaf(w[0][1], y, w[0][0] * x));
        r.y = (a.y + b.y) + fmaf(w[1][2], z, fmaf(w[1][1], y, w[1][0] * x));
        r.z = (a.z + b.z) + fmaf(w[2][2], z, fmaf(w[2][1], y, w[2][0] * x));
        r.w = (a.w + b.w) + fmaf(w[3][2], z, fmaf(w[3][1], y, w[3][0] * x));
        *reinterpret_cast<float4 *>(o) = r;
    };
    const int head = len & 3;
    if (head) {
        int64_t n0 = np[0], n1 = n0, n2 = n0;
        if (head > 1) n1 = np[1];
        if (head > 2) n2 = np[2];
        const float4 b0 = __ldg(reinterpret_cast<const float4 *>(pb + n0 * ld + col));
        float4 b1 = b0, b2 = b0;
        if (head > 1) b1 = __ldg(reinterpret_cast<const float4 *>(pb + n1 * ld + col));
        if (head > 2) b2 = __ldg(reinterpret_cast<const float4 *>(pb + n2 * ld + col));
        emit(op, b0, __ldg(sp), __ldg(sp + 1), __ldg(sp + 2));
        if (head > 1) emit(op + H, b1, __ldg(sp + 3), __ldg(sp + 4), __ldg(sp + 5));
        if (head > 2) emit(op + 2 * H, b2, __ldg(sp + 6), __ldg(sp + 7), __ldg(sp + 8));
        op += head * H;
        sp += head * 3;
    }
#pragma unroll 2
    for (int r = head; r < len; r += 4) {
        const int64_t n0 = np[r], n1 = np[r + 1], n2 = np[r + 2], n3 = np[r + 3];
        const float4 b0 = __ldg(reinterpret_cast<const float4 *>(pb + n0 * ld + col));
        const float4 b1 = __ldg(reinterpret_cast<const float4 *>(pb + n1 * ld + col));
        const float4 b2 = __ldg(reinterpret_cast<const float4 *>(pb + n2 * ld + col));
        const float4 b3 = __ldg(reinterpret_cast<const float4 *>(pb + n3 * ld + col));
        float x[12];
#pragma unroll
        for (int k = 0; k < 12; ++k) x[k] = __ldg(sp + k);
        emit(op, b0, x[0], x[1], x[2]);
        emit(op + H, b1, x[3], x[4], x[5]);
        emit(op + 2 * H, b2, x[6], x[7], x[8]);
        emit(op + 3 * H, b3, x[9], x[10], x[11]);
        op += 4 * H;
        sp += 12;
    }
}

// Each CTA owns R = blockDim / (H/4) row slots; a thread keeps 4 channels x 3 coordinates.
__global__ void __launch_bounds__(kBlock)
k_edge_xyz_wgrad(const float4 *__restrict__ du, const float *__restrict__ sxyz, int64_t rows, int H4, int R,
                 float *__restrict__ partial) {
    extern __shared__ float red[];  // [R][H*3]
    const int lane = threadIdx.x % H4, r = threadIdx.x / H4;
    float acc[4][3];
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[j][0] = acc[j][1] = acc[j][2] = 0.f;
    if (r < R) {
#pragma unroll 4
        for (int64_t e = (int64_t)blockIdx.x * R + r; e < rows; e += (int64_t)gridDim.x * R) {
            const float4 g = __ldcs(du + e * H4 + lane);
            const float *p = sxyz + e * 3;
            const float x = __ldg(p), y = __ldg(p + 1), z = __ldg(p + 2);
            acc[0][0] = fmaf(g.x, x, acc[0][0]); acc[0][1] = fmaf(g.x, y, acc[0][1]); acc[0][2] = fmaf(g.x, z, acc[0][2]);
            acc[1][0] = fmaf(g.y, x, acc[1][0]); acc[1][1] = fmaf(g.y, y, acc[1][1]); acc[1][2] = fmaf(g.y, z, acc[1][2]);
            acc[2][0] = fmaf(g.z, x, acc[2][0]); acc[2][1] = fmaf(g.z, y, acc[2][1]); acc[2][2] = fmaf(g.z, z, acc[2][2]);
            acc[3][0] = fmaf(g.w, x, acc[3][0]); acc[3][1] = fmaf(g.w, y, acc[3][1]); acc[3][2] = fmaf(g.w, z, acc[3][2]);
        }
        float *o = red + (size_t)r * H4 * 12 + lane * 12;
#pragma unroll
        for (int j = 0; j < 4; ++j) { o[j * 3] = acc[j][0]; o[j * 3 + 1] = acc[j][1]; o[j * 3 + 2] = acc[j][2]; }
    }
    __syncthreads();
    const int n = H4 * 12;  // = H*3, laid out [h][c]
    for (int t = threadIdx.x; t < n; t += blockDim.x) {
        float s = 0.f;
        for (int k = 0; k < R; ++k) s += red[(size_t)k * n + t];  // fixed order: deterministic
        partial[(size_t)blockIdx.x * n + t] = s;
    }
}

// Weight gradient of  pool = segmented-max over each centre's rows of (X Wo^T + ...):  only the arg-max
// row of every (centre, channel) carries gradient, so
//     dWo[f,:] = sum_i dpool[i,f] * X[bgs[i] + max_idxs[i,f], :]
// is pn*O row FMAs instead of the csum*O of the dense product dy^T X (k-bar times fewer), and reads
// each X row through L1/L2 (the <= k-bar distinct rows of a centre are shared by its O channels).
// Thread (s, f): centre slot s of the CTA, output channel f, J accumulators.
template <int J>
__global__ void __launch_bounds__(256)
k_pool_linear_wgrad(const float *__restrict__ dpool, const int32_t *__restrict__ midx, const int32_t *__restrict__ lens,
                    const int32_t *__restrict__ bgs, const float *__restrict__ x, int64_t pn, int O, int S,
                    float *__restrict__ partial) {
    extern __shared__ float red[];  // [J][O+1]
    const int f = threadIdx.x % O, s = threadIdx.x / O;
    float acc[J];
#pragma unroll
    for (int j = 0; j < J; ++j) acc[j] = 0.f;
#pragma unroll 2
    for (int64_t i = (int64_t)blockIdx.x * S + s; i < pn; i += (int64_t)gridDim.x * S) {
        const bool live = __ldg(lens + i) > 0;
        const float g = live ? __ldg(dpool + i * O + f) : 0.f;
        const int64_t row = live ? (int64_t)__ldg(bgs + i) + __ldg(midx + i * O + f) : 0;
        const float4 *xr = reinterpret_cast<const float4 *>(x + row * J);
#pragma unroll
        for (int j = 0; j < J; j += 4) {
            const float4 v = __ldg(xr + j / 4);
            acc[j] = fmaf(g, v.x, acc[j]); acc[j + 1] = fmaf(g, v.y, acc[j + 1]);
            acc[j + 2] = fmaf(g, v.z, acc[j + 2]); acc[j + 3] = fmaf(g, v.w, acc[j + 3]);
        }
    }
    for (int round = 0; round < S; ++round) {  // slot 0 first, then 1, ...: a fixed summation order
        if (s == round) {
#pragma unroll
            for (int j = 0; j < J; ++j) {
                if (round == 0) red[j * (O + 1) + f] = acc[j];
                else red[j * (O + 1) + f] += acc[j];
            }
        }
        __syncthreads();
    }
    for (int t = threadIdx.x; t < O * J; t += blockDim.x) {
        const int ff = t / J, j = t - ff * J;
        partial[(size_t)blockIdx.x * O * J + t] = red[j * (O + 1) + ff];
    }
}

// ---- the same weight gradient with the rows staged in shared memory ---------------------------------------------
// The thread-per-(centre, channel) kernel above gathers one x row per thread: 32 different rows per warp load, every row
// read ~O/k-bar times through L1 (2.1 GB of L1 traffic for 0.66 GB of rows at level 0): 2.3-2.5 TB/s of its compulsory
// bytes.  Here a warp walks a contiguous range of centres, i.e. ONE contiguous stream of x rows: the rows of the next
// centre (<= 32 at a time) travel into a per-warp shared-memory tile with 4-byte cp.async (no registers, coalesced, every
// row read once) — 8 bytes at a time — while the current centre is summed: lane = channel f reads ITS arg-max row from the
// tile (row pitch J + 2 words: rows m and m + 16 share banks, equal rows broadcast) and does the same J FMAs as before.
// (A dense dy^T x on mma.sync, 3xTF32, was tried first: k-bar times the flops of this sparse sum, 1.5-2x slower than the
// gather kernel; L2 prefetch of the three streams ahead of the warp: no gain, the kernel is bound by its ~150 instructions
// per centre.)  Fixed centre ranges, fixed fold order: deterministic.
constexpr int kPsWarps = 8, kPsRows = 32;
template <int O, int J>
__global__ void __launch_bounds__(kPsWarps * 32)
k_pool_linear_wgrad_staged(const float *__restrict__ dpool, const int32_t *__restrict__ midx, const int32_t *__restrict__ lens,
                           const int32_t *__restrict__ bgs, const float *__restrict__ x, int64_t pn, float *__restrict__ partial) {
    constexpr int NF = (O + 31) / 32, P = J + 2, kTile = kPsRows * P;
    extern __shared__ __align__(16) float smem_ps[];  // per warp two tiles [kPsRows][J + 2]; afterwards red[O][J + 2]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *tiles = smem_ps + warp * 2 * kTile;
    float acc[NF][J];
#pragma unroll
    for (int q = 0; q < NF; ++q)
#pragma unroll
        for (int j = 0; j < J; ++j) acc[q][j] = 0.f;
    const int64_t W = (int64_t)gridDim.x * kPsWarps, w = (int64_t)blockIdx.x * kPsWarps + warp;
    const int64_t cpw = (pn + W - 1) / W;
    const int64_t i0 = w * cpw, i1 = i0 + cpw < pn ? i0 + cpw : pn;
    struct Meta {
        int len;
        int64_t bg;
        float g[NF];
        int m[NF];
    };
    auto load_meta = [&](int64_t i, Meta &c) {
        c.len = __ldg(lens + i);
        c.bg = __ldg(bgs + i);
#pragma unroll
        for (int q = 0; q < NF; ++q) {
            const int f = 32 * q + lane;
            c.g[q] = f < O ? __ldg(dpool + i * O + f) : 0.f;
            c.m[q] = f < O ? __ldg(midx + i * O + f) : -1;
        }
    };
    // rows [c0, min(len, c0 + kPsRows)) of the centre -> tile (contiguous in x: element e of the piece is x[first + e])
    auto issue = [&](const Meta &c, int c0, float *tile) {
        const int nrows = c.len - c0 < kPsRows ? c.len - c0 : kPsRows;
        const float *src = x + (c.bg + c0) * J;
        // 8 bytes per copy (x rows are 16-byte aligned, the tile rows 8-byte): lane -> (row e / (J/2), column pair e % (J/2))
        constexpr int H = J / 2;
        for (int e = lane; e < nrows * H; e += 32) {
            const int r = e / H, c2 = e - r * H;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((uint32_t)__cvta_generic_to_shared(tile + r * P + 2 * c2)), "l"(src + 2 * e) : "memory");
        }
    };
    if (i0 < i1) {
        Meta cur, nxt;
        load_meta(i0, cur);
        nxt = cur;
        if (i0 + 1 < i1) load_meta(i0 + 1, nxt);
        issue(cur, 0, tiles);
        asm volatile("cp.async.commit_group;" ::: "memory");
        int64_t i = i0;
        int c0 = 0;
        uint32_t u = 0;
        while (i < i1) {
            // the unit after (i, c0): the centre's next 32 rows, or the next centre
            const bool same = c0 + kPsRows < cur.len;
            const int64_t ni = same ? i : i + 1;
            const int nc0 = same ? c0 + kPsRows : 0;
            if (ni < i1) issue(same ? cur : nxt, nc0, tiles + ((u + 1) & 1) * kTile);
            asm volatile("cp.async.commit_group;" ::: "memory");
            asm volatile("cp.async.wait_group 1;" ::: "memory");  // this unit's rows have landed (this lane's copies)
            __syncwarp();
            const float *tile = tiles + (u & 1) * kTile;
#pragma unroll
            for (int q = 0; q < NF; ++q) {
                const int r = cur.m[q] - c0;
                if (r >= 0 && r < kPsRows && cur.m[q] < cur.len) {
                    const float g = cur.g[q];
                    const float2 *row = reinterpret_cast<const float2 *>(tile + r * P);
#pragma unroll
                    for (int j = 0; j < J / 2; ++j) {
                        const float2 v = row[j];
                        acc[q][2 * j] = fmaf(g, v.x, acc[q][2 * j]);
                        acc[q][2 * j + 1] = fmaf(g, v.y, acc[q][2 * j + 1]);
                    }
                }
            }
            __syncwarp();  // the tile is free for the unit after the next
            if (!same) {
                cur = nxt;
                if (ni + 1 < i1) load_meta(ni + 1, nxt);
            }
            i = ni;
            c0 = nc0;
            ++u;
        }
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    // fold the CTA's warps in a fixed order, then one partial per CTA: partial[blockIdx][f][j]
    float *red = smem_ps;
    for (int round = 0; round < kPsWarps; ++round) {
        if (warp == round) {
#pragma unroll
            for (int q = 0; q < NF; ++q) {
                const int f = 32 * q + lane;
                if (f < O) {
#pragma unroll
                    for (int j = 0; j < J; ++j) {
                        if (round == 0) red[f * P + j] = acc[q][j];
                        else red[f * P + j] += acc[q][j];
                    }
                }
            }
        }
        __syncthreads();
    }
    for (int idx = threadIdx.x; idx < O * J; idx += kPsWarps * 32) {
        const int f = idx / J, j = idx - f * J;
        partial[(size_t)blockIdx.x * O * J + idx] = red[f * P + j];
    }
}
}  // namespace

extern "C" int pcs_edge_combine_forward(const float *pa, const float *pb, const float *sxyz, const float *wx,
                                        const int32_t *cidxs, const int32_t *nidxs, const int32_t *nidxs_lens,
                                        const int32_t *nidxs_bgs, int64_t pn, int64_t csum, int64_t H, int64_t ldp,
                                        float *u, pcs_stream_t stream) {
    PCS_REQUIRE(csum >= 0 && pn >= 0 && H > 0, "edge_combine: bad sizes pn=%lld csum=%lld H=%lld", (long long)pn,
                (long long)csum, (long long)H);
    PCS_REQUIRE(H % 4 == 0 && H <= 4 * kMaxH4, "edge_combine: H=%lld must be a multiple of 4, <= %d", (long long)H,
                4 * kMaxH4);
    if (csum == 0) return 0;
    PCS_REQUIRE(pa && pb && sxyz && wx && nidxs && u, "edge_combine: null pointer");
    PCS_REQUIRE(ldp >= H && ldp % 4 == 0, "edge_combine: row stride ldp=%lld must be >= H and a multiple of 4",
                (long long)ldp);
    PCS_REQUIRE(aligned(pa, 16) && aligned(pb, 16) && aligned(u, 16), "edge_combine: pa/pb/u must be 16-byte aligned");
    cudaStream_t s = as_stream(stream);
    if (nidxs_lens && nidxs_bgs && pn > 0 && H >= 16) {  // centre-walking kernel
        RowTiling t = make_tiling(H, true, true);
        const int64_t grid = grid_for(pn, t, kBlock);
        k_edge_combine_walk<<<(unsigned)grid, kBlock, 0, s>>>(pa, pb, sxyz, wx, nidxs, nidxs_lens, nidxs_bgs, pn, H, ldp,
                                                               t.gshift, t.CH, u);
        return check_launch("edge_combine_forward");
    }
    PCS_REQUIRE(cidxs, "edge_combine: cidxs required without (nidxs_lens, nidxs_bgs)");
    const int64_t pieces = csum * (H / 4);
    const int64_t grid = (pieces + kBlock * kU - 1) / (kBlock * kU);
    PCS_REQUIRE(grid < ((int64_t)1 << 31), "edge_combine: tensor too large");
    auto a4 = reinterpret_cast<const float4 *>(pa), b4 = reinterpret_cast<const float4 *>(pb);
    auto u4 = reinterpret_cast<float4 *>(u);
    if (pieces < ((int64_t)1 << 31))
        k_edge_combine<uint32_t><<<(unsigned)grid, kBlock, 0, s>>>(a4, b4, sxyz, wx, cidxs, nidxs, (uint32_t)pieces,
                                                                    (uint32_t)(H / 4), (uint32_t)(ldp / 4), u4);
    else
        k_edge_combine<int64_t><<<(unsigned)grid, kBlock, 0, s>>>(a4, b4, sxyz, wx, cidxs, nidxs, pieces, H / 4, ldp / 4,
                                                                   u4);
    return check_launch("edge_combine_forward");
}

extern "C" int64_t pcs_edge_xyz_wgrad_partials(int64_t csum, int64_t H) {
    if (csum <= 0 || H <= 0 || H % 4 != 0 || H > 4 * kMaxH4) return 0;
    const int64_t R = kBlock / (H / 4);
    const int64_t want = (csum + R * 8 - 1) / (R * 8);  // >= 8 rows per slot
    const int64_t cap = 148 * 6;  // 40 registers x 256 threads: 6 CTAs per SM (4 left the stream at 66 % of the copy rate)
    return want < cap ? want : cap;
}

extern "C" int pcs_edge_xyz_wgrad(const float *du, const float *sxyz, int64_t csum, int64_t H, float *partial,
                                  pcs_stream_t stream) {
    PCS_REQUIRE(csum >= 0 && H > 0, "edge_xyz_wgrad: bad sizes csum=%lld H=%lld", (long long)csum, (long long)H);
    PCS_REQUIRE(H % 4 == 0 && H <= 4 * kMaxH4, "edge_xyz_wgrad: H=%lld must be a multiple of 4, <= %d", (long long)H,
                4 * kMaxH4);
    if (csum == 0) return 0;
    PCS_REQUIRE(du && sxyz && partial, "edge_xyz_wgrad: null pointer");
    PCS_REQUIRE(aligned(du, 16), "edge_xyz_wgrad: du must be 16-byte aligned");
    const int H4 = (int)(H / 4), R = kBlock / H4;
    const int64_t grid = pcs_edge_xyz_wgrad_partials(csum, H);
    const size_t smem = (size_t)R * H * 3 * sizeof(float);
    k_edge_xyz_wgrad<<<(unsigned)grid, kBlock, smem, as_stream(stream)>>>(reinterpret_cast<const float4 *>(du), sxyz,
                                                                         csum, H4, R, partial);
    return check_launch("edge_xyz_wgrad");
}

static inline int pool_wgrad_slots(int64_t O) { return O >= 256 ? 1 : (int)(256 / O); }

extern "C" int64_t pcs_pool_linear_wgrad_partials(int64_t pn, int64_t O) {
    if (pn <= 0 || O <= 0 || O > 256) return 0;
    const int64_t S = pool_wgrad_slots(O);
    const int64_t want = (pn + S * 4 - 1) / (S * 4);  // >= 4 centres per slot
    // CTAs per SM: the staged form holds two row tiles per warp in shared memory (8 warps: 70 KB at J = 32, 102 KB at 48, 135 KB at 64)
    const int64_t cap = 148 * (!opt::pool_wgrad_staged.load() ? 3 : O == 32 ? 3 : O == 48 ? 2 : O == 64 ? 1 : 3);
    return want < cap ? want : cap;
}

extern "C" int pcs_pool_linear_wgrad(const float *dpool, const int32_t *max_idxs, const int32_t *nidxs_lens,
                                     const int32_t *nidxs_bgs, const float *x, int64_t pn, int64_t csum, int64_t O,
                                     int64_t J, float *partial, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && csum >= 0, "pool_linear_wgrad: negative size");
    PCS_REQUIRE(O > 0 && O <= 256, "pool_linear_wgrad: O=%lld out of range (1..256)", (long long)O);
    PCS_REQUIRE(J == 32 || J == 48 || J == 64, "pool_linear_wgrad: row width %lld not built (32, 48, 64)", (long long)J);
    if (pn == 0) return 0;
    PCS_REQUIRE(dpool && max_idxs && nidxs_lens && nidxs_bgs && partial, "pool_linear_wgrad: null pointer");
    PCS_REQUIRE(csum > 0 && x, "pool_linear_wgrad: no rows to gather from");
    PCS_REQUIRE(aligned(x, 16), "pool_linear_wgrad: x must be 16-byte aligned");
    const int S = pool_wgrad_slots(O);
    const int64_t grid = pcs_pool_linear_wgrad_partials(pn, O);
    const size_t smem = (size_t)J * (O + 1) * sizeof(float);
    cudaStream_t s = as_stream(stream);
    if (opt::pool_wgrad_staged.load() && aligned(x, 16)) {  // rows staged in shared memory (same partial layout)
        auto launch = [&](auto kern, int o, int j) -> int {
            const size_t need = (size_t)kPsWarps * 2 * kPsRows * (j + 2) * sizeof(float);
            const size_t red = (size_t)o * (j + 2) * sizeof(float);
            const size_t bytes = need > red ? need : red;
            PCS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
            kern<<<(unsigned)grid, kPsWarps * 32, bytes, s>>>(dpool, max_idxs, nidxs_lens, nidxs_bgs, x, pn, partial);
            return check_launch("pool_linear_wgrad");
        };
        if (O == 32 && J == 32) return launch(k_pool_linear_wgrad_staged<32, 32>, 32, 32);
        if (O == 48 && J == 48) return launch(k_pool_linear_wgrad_staged<48, 48>, 48, 48);
        if (O == 64 && J == 64) return launch(k_pool_linear_wgrad_staged<64, 64>, 64, 64);
    }
    const unsigned block = (unsigned)(O * S);
    if (J == 32) k_pool_linear_wgrad<32><<<(unsigned)grid, block, smem, s>>>(dpool, max_idxs, nidxs_lens, nidxs_bgs, x, pn, (int)O, S, partial);
    else if (J == 48) k_pool_linear_wgrad<48><<<(unsigned)grid, block, smem, s>>>(dpool, max_idxs, nidxs_lens, nidxs_bgs, x, pn, (int)O, S, partial);
    else k_pool_linear_wgrad<64><<<(unsigned)grid, block, smem, s>>>(dpool, max_idxs, nidxs_lens, nidxs_bgs, x, pn, (int)O, S, partial);
    return check_launch("pool_linear_wgrad");
}

// column_sum.cu — out[c] = sum_r x[r, c] for a tall row-major matrix (rows ~ 10^5..10^6, C <= a few hundred):
// the bias gradient of every point-level linear layer around the neighbour ops, and the fold of the
// per-CTA weight-gradient partials the fused kernels emit.  The stock reduction spends 100-150 us on a
// 68 MB [133399, 128] tensor (profiles/r2a_torchprof.log); this is a plain HBM stream.
//
// grid = (row slabs, column chunks of <= 32 vector lanes).  Thread t of a CTA -> column lane t % CW, row
// phase t / CW (exact tiling, the column of a thread never changes); a warp reads up to 512 contiguous
// bytes of a row, four independent loads in flight per thread, per-thread accumulation in row order, then
// a fixed-order shared-memory fold over the row phases.  One partial row per slab; the same
// kernel with a single slab folds the partials.  No atomics: bit-reproducible.
#include "common.cuh"

using namespace pcs;

namespace {
constexpr int kBlock = 256;
constexpr int kSlabCap = 148 * 4;
constexpr int kLanes = 32;  // column lanes per CTA: a warp reads 512 contiguous bytes of a row, 8 row phases per CTA

template <int V>
__global__ void __launch_bounds__(kBlock)
k_column_sum(const float *__restrict__ x, int64_t rows, int64_t C, int CW, int RP, int64_t rows_per_slab,
             float *__restrict__ out) {
    __shared__ float red[kBlock * V];
    const int t = threadIdx.x;
    const int rp = t / CW, lane = t - rp * CW;             // row phase, column lane inside this column chunk
    const int64_t col = ((int64_t)blockIdx.y * CW + lane) * V;
    const int64_t r_bg = (int64_t)blockIdx.x * rows_per_slab;
    const int64_t r_en = r_bg + rows_per_slab < rows ? r_bg + rows_per_slab : rows;
    Vec<V> acc = vzero<V>();
    if (rp < RP && col < C) {
        const float *p = x + col;
        int64_t r = r_bg + rp;
        const int64_t step = RP;
        for (; r + 3 * step < r_en; r += 4 * step) {
            const Vec<V> a = ld<V>(p + r * C), b = ld<V>(p + (r + step) * C), c = ld<V>(p + (r + 2 * step) * C),
                         d = ld<V>(p + (r + 3 * step) * C);
            vadd<V>(acc, a); vadd<V>(acc, b); vadd<V>(acc, c); vadd<V>(acc, d);
        }
        for (; r < r_en; r += step) vadd<V>(acc, ld<V>(p + r * C));
    }
#pragma unroll
    for (int k = 0; k < V; ++k) red[t * V + k] = acc.v[k];
    __syncthreads();
    if (rp == 0 && col < C) {
        Vec<V> s = acc;
        for (int q = 1; q < RP; ++q)
#pragma unroll
            for (int k = 0; k < V; ++k) s.v[k] += red[(q * CW + lane) * V + k];
        st<V>(out + (int64_t)blockIdx.x * C + col, s);
    }
}

struct Plan {
    int V, CW, RP;
    int64_t chunks, slabs, rows_per_slab;
};
Plan make_plan(int64_t rows, int64_t C, bool al16, bool al8) {
    Plan p;
    p.V = (al16 && C % 4 == 0) ? 4 : ((al8 && C % 2 == 0) ? 2 : 1);
    const int64_t lanes = C / p.V;
    p.CW = (int)(lanes < kLanes ? lanes : kLanes);
    p.RP = kBlock / p.CW;
    p.chunks = (lanes + p.CW - 1) / p.CW;
    // slab count independent of the vector width (the caller sizes `partial` before alignment is known):
    // >= 32 rows per slab, at most kSlabCap slabs (= partial rows)
    int64_t slabs = (rows + 31) / 32;
    if (slabs > kSlabCap) slabs = kSlabCap;
    if (slabs < 1) slabs = 1;
    p.rows_per_slab = (rows + slabs - 1) / slabs;
    p.slabs = (rows + p.rows_per_slab - 1) / p.rows_per_slab;
    return p;
}

int launch(const float *x, int64_t rows, int64_t C, const Plan &p, float *out, cudaStream_t s) {
    dim3 grid((unsigned)p.slabs, (unsigned)p.chunks);
    switch (p.V) {
        case 4: k_column_sum<4><<<grid, kBlock, 0, s>>>(x, rows, C, p.CW, p.RP, p.rows_per_slab, out); break;
        case 2: k_column_sum<2><<<grid, kBlock, 0, s>>>(x, rows, C, p.CW, p.RP, p.rows_per_slab, out); break;
        default: k_column_sum<1><<<grid, kBlock, 0, s>>>(x, rows, C, p.CW, p.RP, p.rows_per_slab, out); break;
    }
    return check_launch("column_sum");
}
}  // namespace

extern "C" int64_t pcs_column_sum_partials(int64_t rows, int64_t C) {
    if (rows <= 0 || C <= 0) return 0;
    const Plan p = make_plan(rows, C, true, true);
    return p.slabs > 1 ? p.slabs : 0;
}

extern "C" int pcs_column_sum(const float *x, int64_t rows, int64_t C, float *partial, float *out,
                              pcs_stream_t stream) {
    PCS_REQUIRE(rows >= 0 && C >= 0, "column_sum: negative size");
    if (C == 0) return 0;
    PCS_REQUIRE(out, "column_sum: out required");
    cudaStream_t s = as_stream(stream);
    if (rows == 0) {
        PCS_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * C, s));
        return 0;
    }
    PCS_REQUIRE(x, "column_sum: x required");
    const Plan p = make_plan(rows, C, aligned(x, 16), aligned(x, 8));
    if (p.slabs == 1) return launch(x, rows, C, make_plan(rows, C, aligned(x, 16) && aligned(out, 16),
                                                          aligned(x, 8) && aligned(out, 8)), out, s);
    PCS_REQUIRE(partial, "column_sum: partial [pcs_column_sum_partials(rows, C), C] required");
    Plan p1 = make_plan(rows, C, aligned(x, 16) && aligned(partial, 16), aligned(x, 8) && aligned(partial, 8));
    if (int rc = launch(x, rows, C, p1, partial, s)) return rc;
    Plan p2 = make_plan(p1.slabs, C, aligned(partial, 16) && aligned(out, 16), aligned(partial, 8) && aligned(out, 8));
    p2.slabs = 1;  // fold all partial rows in one CTA per column chunk
    p2.rows_per_slab = p1.slabs;
    return launch(partial, p1.slabs, C, p2, out, s);
}

// pooling_index.cu — a13/a14 ComputePermutationInfo[WithClass], a17 ComputeRepermutationInfo, the
// transpose CSR used by the deterministic NeighborGather, and the int32 exclusive scan helper.
// CUB device primitives (radix sort, scan — header templates shipped with the toolkit, compiled
// into this library) stand where the reference calls thrust (ComputePermutationInfo.cu:117-144).
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

using namespace pcs;

namespace {
constexpr int kBlock = 256;

// key = ((x<<16)+y<<16)+z [<<16 + class] in signed 64-bit arithmetic, exactly the sequence of
// computeWholeIdxs[WithClass]Kernel (ComputePermutationInfo.cu:10-44).
// block_ids (extension, no classes): the 48-bit key must be in range and the block id is added
// above bit 48.  flags (vn_dev[1]): bit0 = a key does not fit `sort_bits` (caller re-runs with
// 64), bit1 = key out of the 48-bit range while batching (unsupported input).
__global__ void __launch_bounds__(kBlock)
k_build_keys(const int32_t *__restrict__ vidx, const int32_t *__restrict__ classes,
             const int32_t *__restrict__ block_ids, int64_t pn, int sort_bits, int64_t *__restrict__ keys,
             int32_t *__restrict__ vals, int32_t *__restrict__ flags) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pn) return;
    int64_t k = vidx[i * 3];
    k <<= 16; k += vidx[i * 3 + 1];
    k <<= 16; k += vidx[i * 3 + 2];
    int bad = 0;
    if (classes) {
        k <<= 16; k += classes[i];
    } else if (block_ids) {
        if (k < 0 || (k >> 48) != 0) bad = 2;
        k += (int64_t)block_ids[i] << 48;
    } else if (sort_bits < 64) {
        if (k < 0 || (k >> sort_bits) != 0) bad = 1;
    }
    keys[i] = k;
    vals[i] = (int32_t)i;
    if (bad) atomicOr(flags, bad);
}

// After the sort: p2o, head flags (1 at the first point of every voxel run).
__global__ void __launch_bounds__(kBlock)
k_heads_inverse(const int64_t *__restrict__ skeys, const int32_t *__restrict__ o2p, int64_t pn,
                int32_t *__restrict__ p2o, int32_t *__restrict__ heads) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pn) return;
    p2o[o2p[p]] = (int32_t)p;
    heads[p] = (p == 0 || skeys[p] != skeys[p - 1]) ? 1 : 0;
}

// cens_incl = inclusive scan of heads.  cens[p] = cens_incl[p]-1; a head writes begs[v] = p; the
// last point writes vn.  Then lens[v] = begs[v+1]-begs[v] (second kernel, begs[vn] = pn implied).
__global__ void __launch_bounds__(kBlock)
k_finish_cens(const int32_t *__restrict__ heads, int32_t *__restrict__ cens, int64_t pn,
              int32_t *__restrict__ begs, int32_t *__restrict__ vn_dev) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pn) return;
    const int v = cens[p] - 1;
    cens[p] = v;
    if (heads[p]) begs[v] = (int32_t)p;
    if (p == pn - 1) vn_dev[0] = v + 1;
}
__global__ void __launch_bounds__(kBlock)
k_lens_from_begs(const int32_t *__restrict__ begs, const int32_t *__restrict__ vn_dev, int64_t pn,
                 int32_t *__restrict__ lens) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int vn = vn_dev[0];
    if (v >= vn) return;
    const int next = (v + 1 < vn) ? begs[v + 1] : (int32_t)pn;
    lens[v] = next - begs[v];
}

// ---- a17 (ComputeRepermutationInfo.cu:20-88) --------------------------------------------------
__global__ void __launch_bounds__(kBlock)
k_reper_lens(const int32_t *__restrict__ o2p2, const int32_t *__restrict__ lens, int64_t pn2,
             int32_t *__restrict__ rlens) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < pn2) rlens[i] = lens[o2p2[i]];
}
// 8 lanes per level-2 position fill its run: o2p1[rbeg+j] = obeg+j, p2o1[obeg+j] = rbeg+j, cens = i.
__global__ void __launch_bounds__(kBlock)
k_reper_fill(const int32_t *__restrict__ o2p2, const int32_t *__restrict__ begs,
             const int32_t *__restrict__ rlens, const int32_t *__restrict__ rbegs, int64_t pn2,
             int32_t *__restrict__ o2p1, int32_t *__restrict__ p2o1, int32_t *__restrict__ rcens) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t >> 3;
    if (i >= pn2) return;
    const int len = rlens[i], rbeg = rbegs[i], obeg = begs[o2p2[i]];
    for (int j = (int)(t & 7); j < len; j += 8) {
        o2p1[rbeg + j] = obeg + j;
        p2o1[obeg + j] = rbeg + j;
        rcens[rbeg + j] = (int32_t)i;
    }
}

// ---- transpose CSR ----------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock)
k_iota(int32_t *__restrict__ v, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = (int32_t)i;
}
// skeys sorted ascending: tbgs[n] = first position p with skeys[p] >= n, tbgs[pn] = csum.
__global__ void __launch_bounds__(kBlock)
k_row_starts(const int32_t *__restrict__ skeys, int64_t csum, int64_t pn, int32_t *__restrict__ tbgs) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p > csum) return;
    const int64_t lo = (p == 0) ? 0 : (int64_t)skeys[p - 1] + 1;
    const int64_t hi = (p == csum) ? pn : (int64_t)skeys[p];
    for (int64_t n = lo; n <= hi && n <= pn; ++n) tbgs[n] = (int32_t)p;
}

struct Carve {  // bump allocator over the caller's workspace
    char *base; size_t off, cap;
    template <typename T> T *take(size_t n) {
        off = align_up(off, 256);
        T *p = reinterpret_cast<T *>(base + off);
        off += n * sizeof(T);
        return p;
    }
};
}  // namespace

// ---- exclusive scan ------------------------------------------------------------------------------
__global__ void k_scan_total(const int32_t *in, const int32_t *out, int64_t n, int32_t *total) {
    total[0] = (n > 0) ? in[n - 1] + out[n - 1] : 0;  // the reference's count1+count2, SearchNeighborhood.cu:148-152
}

extern "C" size_t pcs_exclusive_scan_workspace_bytes(int64_t n) {
    size_t b = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, b, (const int32_t *)nullptr, (int32_t *)nullptr, (int)(n > 0 ? n : 1));
    return align_up(b, 256) + 256;
}
extern "C" int pcs_exclusive_scan_i32(const int32_t *in, int64_t n, int32_t *out, int32_t *total_dev,
                                      void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(n >= 0 && n < (int64_t)1 << 31, "exclusive_scan: bad n");
    cudaStream_t s = as_stream(stream);
    if (n > 0) {
        size_t need = 0;
        cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, (int)n, s);
        PCS_REQUIRE(workspace_bytes >= need, "exclusive_scan: workspace %zu < %zu", workspace_bytes, need);
        PCS_CUDA(cub::DeviceScan::ExclusiveSum(workspace, need, in, out, (int)n, s));
        g_launches.fetch_add(2);
    }
    if (total_dev) { k_scan_total<<<1, 1, 0, s>>>(in, out, n, total_dev); return check_launch("scan_total"); }
    return 0;
}

// ---- a13 / a14 -------------------------------------------------------------------------------
static size_t perm_sort_temp_bytes(int64_t pn) {
    size_t b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, b, (const int64_t *)nullptr, (int64_t *)nullptr,
                                    (const int32_t *)nullptr, (int32_t *)nullptr, (int)(pn > 0 ? pn : 1));
    return b;
}
static size_t scan_temp_bytes(int64_t n) {
    size_t b = 0;
    cub::DeviceScan::InclusiveSum(nullptr, b, (const int32_t *)nullptr, (int32_t *)nullptr, (int)(n > 0 ? n : 1));
    return b;
}
extern "C" size_t pcs_compute_permutation_info_workspace_bytes(int64_t pn) {
    if (pn < 1) pn = 1;
    // keys, sorted keys (int64), iota vals, heads (int32), cub temp
    return 2 * align_up(pn * 8, 256) + 2 * align_up(pn * 4, 256) +
           align_up(perm_sort_temp_bytes(pn) > scan_temp_bytes(pn) ? perm_sort_temp_bytes(pn) : scan_temp_bytes(pn), 256) + 2048;
}

// sort_bits is carried in the low byte of workspace_bytes' caller contract?  No: it is chosen here —
// 48 unless classes/block_ids are present; a set bit0 in vn_dev[1] tells the caller to call
// pcs_compute_permutation_info64 (below) instead.
static int permutation_info_impl(const int32_t *voxel_idxs, const int32_t *classes, const int32_t *block_ids,
                                 int64_t pn, int sort_bits, int32_t *o2p, int32_t *p2o, int32_t *lens,
                                 int32_t *begs, int32_t *cens, int32_t *vn_dev, void *workspace,
                                 size_t workspace_bytes, cudaStream_t s) {
    PCS_REQUIRE(pn >= 0 && pn < ((int64_t)1 << 31), "compute_permutation_info: bad pn");
    PCS_REQUIRE(!(classes && block_ids), "compute_permutation_info: classes and block_ids together are unsupported");
    PCS_CUDA(cudaMemsetAsync(vn_dev, 0, 2 * sizeof(int32_t), s));
    if (pn == 0) return 0;
    PCS_REQUIRE(workspace_bytes >= pcs_compute_permutation_info_workspace_bytes(pn),
                "compute_permutation_info: workspace too small");
    Carve c{static_cast<char *>(workspace), 0, workspace_bytes};
    int64_t *keys = c.take<int64_t>(pn);
    int64_t *skeys = c.take<int64_t>(pn);
    int32_t *vals = c.take<int32_t>(pn);
    int32_t *heads = c.take<int32_t>(pn);
    size_t tb = perm_sort_temp_bytes(pn), sb = scan_temp_bytes(pn);
    size_t temp_bytes = tb > sb ? tb : sb;
    void *temp = c.take<char>(temp_bytes);
    const int64_t grid = (pn + kBlock - 1) / kBlock;
    if (classes || block_ids) sort_bits = 64;
    k_build_keys<<<grid, kBlock, 0, s>>>(voxel_idxs, classes, block_ids, pn, sort_bits, keys, vals, vn_dev + 1);
    if (check_launch("build_keys")) return 1;
    // signed 64-bit keys, stable LSD radix sort == thrust::sort_by_key on `long`
    PCS_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys, skeys, vals, o2p, (int)pn, 0, sort_bits, s));
    g_launches.fetch_add(sort_bits / 8 + 1);
    k_heads_inverse<<<grid, kBlock, 0, s>>>(skeys, o2p, pn, p2o, heads);
    if (check_launch("heads_inverse")) return 1;
    PCS_CUDA(cub::DeviceScan::InclusiveSum(temp, temp_bytes, heads, cens, (int)pn, s));
    g_launches.fetch_add(2);
    k_finish_cens<<<grid, kBlock, 0, s>>>(heads, cens, pn, begs, vn_dev);
    if (check_launch("finish_cens")) return 1;
    k_lens_from_begs<<<grid, kBlock, 0, s>>>(begs, vn_dev, pn, lens);
    return check_launch("lens_from_begs");
}

extern "C" int pcs_compute_permutation_info(const int32_t *voxel_idxs, const int32_t *classes,
                                            const int32_t *block_ids, int64_t pn, int32_t *o2p, int32_t *p2o,
                                            int32_t *lens, int32_t *begs, int32_t *cens, int32_t *vn_dev,
                                            void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    return permutation_info_impl(voxel_idxs, classes, block_ids, pn, 48, o2p, p2o, lens, begs, cens, vn_dev,
                                 workspace, workspace_bytes, as_stream(stream));
}
extern "C" int pcs_compute_permutation_info64(const int32_t *voxel_idxs, const int32_t *classes,
                                              const int32_t *block_ids, int64_t pn, int32_t *o2p, int32_t *p2o,
                                              int32_t *lens, int32_t *begs, int32_t *cens, int32_t *vn_dev,
                                              void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    return permutation_info_impl(voxel_idxs, classes, block_ids, pn, 64, o2p, p2o, lens, begs, cens, vn_dev,
                                 workspace, workspace_bytes, as_stream(stream));
}

// ---- a17 ---------------------------------------------------------------------------------------
extern "C" size_t pcs_compute_repermutation_info_workspace_bytes(int64_t pn2) {
    return pcs_exclusive_scan_workspace_bytes(pn2);
}
extern "C" int pcs_compute_repermutation_info(const int32_t *o2p_idxs2, const int32_t *lens, const int32_t *begs,
                                              int64_t pn1, int64_t pn2, int32_t *reper_o2p1, int32_t *reper_p2o1,
                                              int32_t *reper_lens, int32_t *reper_begs, int32_t *reper_cens,
                                              void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn1 >= 0 && pn2 >= 0, "compute_repermutation_info: negative size");
    if (pn2 == 0) return 0;
    cudaStream_t s = as_stream(stream);
    k_reper_lens<<<(pn2 + kBlock - 1) / kBlock, kBlock, 0, s>>>(o2p_idxs2, lens, pn2, reper_lens);
    if (check_launch("reper_lens")) return 1;
    if (pcs_exclusive_scan_i32(reper_lens, pn2, reper_begs, nullptr, workspace, workspace_bytes, stream)) return 1;
    k_reper_fill<<<(pn2 * 8 + kBlock - 1) / kBlock, kBlock, 0, s>>>(o2p_idxs2, begs, reper_lens, reper_begs, pn2,
                                                                     reper_o2p1, reper_p2o1, reper_cens);
    return check_launch("reper_fill");
}

// ---- transpose CSR -------------------------------------------------------------------------------
static size_t tr_sort_temp_bytes(int64_t csum) {
    size_t b = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, b, (const int32_t *)nullptr, (int32_t *)nullptr,
                                    (const int32_t *)nullptr, (int32_t *)nullptr, (int)(csum > 0 ? csum : 1));
    return b;
}
// Transpose of a SYMMETRIC CSR with ascending rows (what the radius / annulus searches return: j in N(i) <=> i in
// N(j), each row ascending in j).  Row n of the transpose lists the edges e with nidxs[e] == n in ascending e,
// i.e. one edge per centre i in N(n), in ascending i: exactly row n of the graph itself.  So tbgs = bgs and
// tperm[bgs[n]+t] = bgs[i] + (rank of n in row i), i = nidxs[bgs[n]+t] — one binary search per edge in a row
// that is nearby in memory, instead of a 3-pass radix sort of all edges.  A miss (graph not symmetric) sets *fail.
__global__ void __launch_bounds__(kBlock)
k_transpose_symmetric(const int32_t *__restrict__ nidxs, const int32_t *__restrict__ lens, const int32_t *__restrict__ bgs,
                      const int32_t *__restrict__ cidxs, int64_t pn, int64_t csum, int32_t *__restrict__ tperm,
                      int32_t *__restrict__ tbgs, int32_t *__restrict__ fail) {
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p <= pn) tbgs[p] = (p == pn) ? (int32_t)csum : __ldg(bgs + p);
    if (p >= csum) return;
    const int32_t n = __ldg(cidxs + p), i = __ldg(nidxs + p);
    const int32_t b = __ldg(bgs + i);
    int lo = 0, hi = __ldg(lens + i);  // first position in row i with value >= n
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(nidxs + b + mid) < n) lo = mid + 1; else hi = mid;
    }
    if (lo >= __ldg(lens + i) || __ldg(nidxs + b + lo) != n) { *fail = 1; tperm[p] = (int32_t)p; return; }
    tperm[p] = b + lo;
}

extern "C" size_t pcs_csr_transpose_workspace_bytes(int64_t pn, int64_t csum) {
    if (csum < 1) csum = 1;
    return 2 * align_up(csum * 4, 256) + align_up(tr_sort_temp_bytes(csum), 256) + 1024;
}
extern "C" int pcs_csr_transpose(const int32_t *nidxs, int64_t pn, int64_t csum, int32_t *tperm, int32_t *tbgs,
                                 void *workspace, size_t workspace_bytes, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && csum >= 0 && csum < ((int64_t)1 << 31), "csr_transpose: bad size");
    cudaStream_t s = as_stream(stream);
    PCS_REQUIRE(workspace_bytes >= pcs_csr_transpose_workspace_bytes(pn, csum), "csr_transpose: workspace too small");
    Carve c{static_cast<char *>(workspace), 0, workspace_bytes};
    int32_t *skeys = c.take<int32_t>(csum > 0 ? csum : 1);
    int32_t *iota = c.take<int32_t>(csum > 0 ? csum : 1);
    size_t temp_bytes = tr_sort_temp_bytes(csum);
    void *temp = c.take<char>(temp_bytes);
    if (csum > 0) {
        k_iota<<<(csum + kBlock - 1) / kBlock, kBlock, 0, s>>>(iota, csum);
        if (check_launch("iota")) return 1;
        int bits = 1;
        while (bits < 31 && ((int64_t)1 << bits) < pn) ++bits;
        PCS_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, nidxs, skeys, iota, tperm, (int)csum, 0, bits, s));
        g_launches.fetch_add((bits + 7) / 8 + 1);
    }
    k_row_starts<<<(csum + 1 + kBlock - 1) / kBlock, kBlock, 0, s>>>(skeys, csum, pn, tbgs);
    return check_launch("row_starts");
}

extern "C" int pcs_csr_transpose_symmetric(const int32_t *nidxs, const int32_t *nidxs_lens, const int32_t *nidxs_bgs,
                                           const int32_t *cidxs, int64_t pn, int64_t csum, int32_t *tperm, int32_t *tbgs,
                                           int32_t *fail_flag, pcs_stream_t stream) {
    PCS_REQUIRE(pn >= 0 && csum >= 0 && csum < ((int64_t)1 << 31), "csr_transpose_symmetric: bad size");
    PCS_REQUIRE(tbgs && fail_flag && (csum == 0 || (nidxs && nidxs_lens && nidxs_bgs && cidxs && tperm)),
                "csr_transpose_symmetric: null pointer");
    const int64_t n = (csum > pn + 1) ? csum : pn + 1;
    k_transpose_symmetric<<<(unsigned)((n + kBlock - 1) / kBlock), kBlock, 0, as_stream(stream)>>>(
        nidxs, nidxs_lens, nidxs_bgs, cidxs, pn, csum, tperm, tbgs, fail_flag);
    return check_launch("csr_transpose_symmetric");
}

// common.cu — error string, launch counter, version.
#include "common.cuh"
#include <string.h>
#include <stdarg.h>

namespace pcs {
thread_local char g_err[512] = "";
std::atomic<int64_t> g_launches{0};

int fail(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

int check_launch(const char *what) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail("%s launch: %s", what, cudaGetErrorString(e));
    return 0;
}
namespace opt {
// pcs_neighbor_scatter: rows at least this wide go to the centre-walking kernel (INT64_MAX = never).
// Defaults from tools/ab_rows.py on B200 (64 x 4096-pt blocks, mean degree 12): walk/flat % of HBM peak
//   copy  C=12: 43/58  16: 67/62  32: 80/67  64: 92/82  128: 97/91;  odd 35: 53/57  131: 60/61
//   diff  C=12: 48/42  16: 70/41  32: 84/43  64: 95/53  128: 98/62;  odd 3: 24/24  35: 56/40  131: 60/47
std::atomic<int64_t> scatter_walk_min_c_diff{12}, scatter_walk_min_c_diff_odd{5}, scatter_walk_min_c_copy{16},
    scatter_walk_min_c_copy_odd{INT64_MAX};
// pcs_dense_mlp3_*: 1 = tensor-core kernels (error-compensated TF32), 0 = fp32 FFMA kernels
std::atomic<int64_t> mlp_tensor_cores{1};
std::atomic<int64_t> pool_wgrad_staged{1};
// walkers (segment_reduce.cu, neighbor_ops.cu): rows in flight per lane when the lane is narrower than 16 bytes
std::atomic<int64_t> walk_rows_narrow{8};
// seg_stream.cu (TMA-staged stream reductions / broadcasts): 0 = never, 1 = where measured faster, 2 = always
std::atomic<int64_t> seg_stream{1};
// flat (output-indexed) kernels for a6 / a8 below this many row bytes
std::atomic<int64_t> flat_small_c{64};
// a11 eliminate_center: 1 = edge-parallel tile kernel, 0 = the sub-warp-per-row ballot kernel
std::atomic<int64_t> elim_tile{1};
}  // namespace opt
}  // namespace pcs

extern "C" int pcs_set_option(const char *name, int64_t value) {
    using namespace pcs;
    PCS_REQUIRE(name != nullptr, "set_option: null name");
    struct { const char *n; std::atomic<int64_t> *v; bool threshold; } opts[] = {
        {"scatter_walk_min_c_diff", &opt::scatter_walk_min_c_diff, true},
        {"scatter_walk_min_c_diff_odd", &opt::scatter_walk_min_c_diff_odd, true},
        {"scatter_walk_min_c_copy", &opt::scatter_walk_min_c_copy, true},
        {"scatter_walk_min_c_copy_odd", &opt::scatter_walk_min_c_copy_odd, true},
        {"mlp_tensor_cores", &opt::mlp_tensor_cores, false},
        {"pool_wgrad_staged", &opt::pool_wgrad_staged, false},
        {"walk_rows_narrow", &opt::walk_rows_narrow, false},
        {"seg_stream", &opt::seg_stream, false},
        {"flat_small_c", &opt::flat_small_c, false},
        {"elim_tile", &opt::elim_tile, false}};
    for (auto &o : opts)
        if (strcmp(o.n, name) == 0) { o.v->store((o.threshold && value < 0) ? INT64_MAX : value); return 0; }
    return fail("set_option: unknown option '%s'", name);
}

extern "C" const char *pcs_last_error(void) { return pcs::g_err; }
extern "C" int pcs_version(void) { return 100; }
extern "C" int64_t pcs_launch_count(void) { return pcs::g_launches.load(); }

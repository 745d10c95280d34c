// scan_static.cu — pcs_exclusive_scan_i32_static: the int32 exclusive prefix sum of pcs_exclusive_scan_i32
// (pooling_index.cu) for callers that cannot hand over a workspace they would have to cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid allocation).  No counterpart in the
// reference.
//
// The scratch CUB's scan needs (its decoupled-look-back tile states) lives in a file-scope device array of fixed
// size: every GPU that loads the library holds kScratchBytes until exit.  Consequences, all documented in the header:
//   * n is limited to PCS_SCAN_STATIC_MAX_N elements (larger sizes are rejected with an error, nothing is launched);
//   * the entry point calls no cudaMalloc / cudaFree / allocating library routine;
//   * every call on a device shares the one array, and it keeps the previous call's contents.  Nothing has to be
//     cleared by hand: the scan's first kernel (CUB's DeviceScanInitKernel) re-initialises every tile state it is
//     going to use before the scan kernel reads any — but two calls that OVERLAP on the GPU (different streams, or
//     two host threads on one device) race on the array.  Calls on one stream are ordered and safe.
#include "common.cuh"
#include <cub/device/device_scan.cuh>
#include <mutex>

using namespace pcs;

namespace {
constexpr int64_t kMaxN = (int64_t)1 << 22;       // 4 194 304 elements (16 MB of int32): PCS_SCAN_STATIC_MAX_N
constexpr size_t kScratchBytes = 256 * 1024;      // >> CUB's need at kMaxN (~25 KB of tile states), checked per call
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_scan_scratch[kScratchBytes];

// In host code `g_scan_scratch` is only a stand-in for the symbol, NOT a device address: the address on a given
// GPU comes from cudaGetSymbolAddress with that GPU current, once per device.
std::mutex g_mu;
void *g_scratch_ptr[kMaxDevices] = {nullptr};

int scratch_for_current_device(void **ptr) {
    int dev = -1;
    PCS_CUDA(cudaGetDevice(&dev));
    PCS_REQUIRE(dev >= 0 && dev < kMaxDevices, "exclusive_scan_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        PCS_CUDA(cudaGetSymbolAddress(&p, g_scan_scratch));
        PCS_REQUIRE(p != nullptr, "exclusive_scan_static: cudaGetSymbolAddress returned null");
        g_scratch_ptr[dev] = p;
    }
    *ptr = g_scratch_ptr[dev];
    return 0;
}

__global__ void k_scan_total_static(const int32_t *__restrict__ in, const int32_t *__restrict__ out, int64_t n,
                                    int32_t *__restrict__ total) {
    total[0] = (n > 0) ? in[n - 1] + out[n - 1] : 0;
}
}  // namespace

extern "C" int64_t pcs_exclusive_scan_static_max_n(void) { return kMaxN; }

extern "C" int pcs_exclusive_scan_i32_static(const int32_t *in, int64_t n, int32_t *out, int32_t *total_dev,
                                             pcs_stream_t stream) {
    PCS_REQUIRE(n >= 0, "exclusive_scan_static: negative n");
    PCS_REQUIRE(n <= kMaxN, "exclusive_scan_static: n = %lld exceeds the static scratch limit of %lld elements "
                            "(use pcs_exclusive_scan_i32 with a caller workspace)", (long long)n, (long long)kMaxN);
    cudaStream_t s = as_stream(stream);
    if (n > 0) {
        void *scratch = nullptr;
        if (scratch_for_current_device(&scratch)) return 1;
        size_t need = 0;
        PCS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, need, in, out, (int)n, s));  // size query: no launch, no allocation
        PCS_REQUIRE(need <= kScratchBytes, "exclusive_scan_static: scan needs %zu bytes of scratch, the static array has %zu",
                    need, kScratchBytes);
        PCS_CUDA(cub::DeviceScan::ExclusiveSum(scratch, need, in, out, (int)n, s));
        g_launches.fetch_add(2);
    }
    if (total_dev) {
        k_scan_total_static<<<1, 1, 0, s>>>(in, out, n, total_dev);
        return check_launch("scan_total_static");
    }
    return 0;
}

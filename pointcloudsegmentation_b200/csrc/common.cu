// common.cu — error string, launch counter, version.
#include "common.cuh"
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
}  // namespace pcs

extern "C" const char *pcs_last_error(void) { return pcs::g_err; }
extern "C" int pcs_version(void) { return 100; }
extern "C" int64_t pcs_launch_count(void) { return pcs::g_launches.load(); }

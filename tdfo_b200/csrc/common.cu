// Error reporting and library-wide bookkeeping for the C ABI.
#include "common.cuh"
#include <stdarg.h>
#include <string.h>

namespace tdfo {

static thread_local char g_error[512] = "";
std::atomic<unsigned long long> g_launch_count{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error in %s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
    return TDFO_ERR_CUDA;
}

}  // namespace tdfo

extern "C" {

int tdfo_abi_version(void) { return TDFO_ABI_VERSION; }

const char* tdfo_last_error(void) { return tdfo::g_error; }

// Number of kernels this library has launched since it was loaded (host counter;
// bench.py reads it around the timed region for its "gpu_launches" claim).
unsigned long long tdfo_launch_count(void) { return tdfo::g_launch_count.load(std::memory_order_relaxed); }

}  // extern "C"

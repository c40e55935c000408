// Error reporting and size helpers of the C ABI.
#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>

#ifndef PDT_PDL_DEFAULT_MASK
// head_bwd (launch overlap only), conv2_bwd_data (weights staged early), conv1_bwd_weight (input band
// staged early): the subset measured faster AND whose pre-wait reads are complete before the step starts
#define PDT_PDL_DEFAULT_MASK 0x1c6
#endif

namespace pdt {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// PDT_PDL is a bit mask over the kernels (bit = the `which` passed to launch_pdl); default: see below.
bool pdl_enabled(int which) {
    static long mask = -2;
    if (mask == -2) {
        const char* e = getenv("PDT_PDL");
        mask = e ? strtol(e, nullptr, 0) : PDT_PDL_DEFAULT_MASK;
    }
    return (mask >> which) & 1;
}

int carveout_percent() {
    static int v = -2;
    if (v == -2) {
        const char* e = getenv("PDT_CARVEOUT");
        v = e ? atoi(e) : -1;
    }
    return v;
}

int check_launch(const char* what) {
    static int sync_debug = -1;     // PDT_SYNC_DEBUG=1: synchronise after every launch so a fault names its kernel
    if (sync_debug < 0) {
        const char* d = getenv("PDT_SYNC_DEBUG");
        sync_debug = (d && d[0] == '1') ? 1 : 0;
    }
    if (sync_debug) cudaDeviceSynchronize();
    const cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return PDT_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return PDT_ECUDA;
}
}  // namespace pdt

extern "C" const char* pdt_last_error(void) { return pdt::g_err; }
extern "C" int pdt_version(void) { return 100; }
extern "C" int64_t pdt_num_tiles(int64_t n) { return pdt::num_tiles(n); }

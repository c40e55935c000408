// Error reporting and size helpers of the C ABI.
#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>

namespace pdt {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

bool pdl_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("PDT_PDL");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

int check_launch(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return PDT_OK;
    set_error("%s: %s", what, cudaGetErrorString(e));
    return PDT_ECUDA;
}
}  // namespace pdt

extern "C" const char* pdt_last_error(void) { return pdt::g_err; }
extern "C" int pdt_version(void) { return 100; }
extern "C" int64_t pdt_num_tiles(int64_t n) { return pdt::num_tiles(n); }

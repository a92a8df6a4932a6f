// api.cu -- error reporting and version of the C ABI (include/isogs_rast.h).
#include <cstdarg>
#include <cstdio>

#include "common.cuh"

namespace isogs {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return ISOGS_E_CUDA;
}

} // namespace isogs

extern "C" const char* isogs_last_error(void) { return isogs::g_err; }
extern "C" int isogs_version(void) { return 100; }

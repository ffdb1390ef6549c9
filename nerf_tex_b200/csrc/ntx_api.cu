// Library-level entry points: version, error string, device info.
#include <stdarg.h>

#include "ntx_common.cuh"

static thread_local char g_err[512] = "";

void ntx_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

NTX_API int ntx_version(void) { return NTX_VERSION; }

NTX_API const char* ntx_last_error(void) { return g_err; }

int ntx_sm_count() {
    static int cached[64];
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

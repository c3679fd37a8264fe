// acx_common.cu -- error text, version, device probe
#include <stdarg.h>
#include <stdio.h>

#include "../../include/acx.h"
#include "acx_common.h"

static thread_local char g_err[512] = "";

int acx_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

extern "C" const char* acx_last_error(void) { return g_err; }
extern "C" int acx_version(void) { return 100; }
extern "C" int acx_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
    return n;
}
extern "C" int acx_set_device(int dev) {
    cudaError_t e = cudaSetDevice(dev);
    if (e != cudaSuccess) return acx_fail(ACX_ERR_CUDA, "cudaSetDevice(%d): %s", dev, cudaGetErrorString(e));
    return ACX_OK;
}

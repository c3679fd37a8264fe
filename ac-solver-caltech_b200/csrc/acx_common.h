// acx_common.h -- host-side error plumbing shared by the translation units of libacx.so
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>

#define ACX_MAX_DEVICES 64

int acx_fail(int code, const char* fmt, ...);

#define ACX_CUDA(expr)                                                                         \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess)                                                                 \
            return acx_fail(-2, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

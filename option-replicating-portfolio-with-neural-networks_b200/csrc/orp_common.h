// Shared host-side helpers for liborp_b200.so (error reporting, launch checks).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstdint>

#include "../../include/orp_b200.h"

namespace orp {

char* last_error_buffer();   // thread-local, 512 bytes

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

inline int cuda_fail(cudaError_t e, const char* what) {
    return fail(ORP_E_CUDA, "%s: %s", what, cudaGetErrorString(e));
}

#define ORP_CUDA(call)                                            \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return orp::cuda_fail(e__, #call); \
    } while (0)

#define ORP_LAUNCH_CHECK(name)                                     \
    do {                                                           \
        cudaError_t e__ = cudaGetLastError();                      \
        if (e__ != cudaSuccess) return orp::cuda_fail(e__, name);  \
    } while (0)

#define ORP_REQUIRE(cond, ...)                                     \
    do {                                                           \
        if (!(cond)) return orp::fail(ORP_E_VALUE, __VA_ARGS__);   \
    } while (0)

int sm_count();   // multiprocessors of the current device (cached per device)

}  // namespace orp

// Shared host-side helpers for the C-ABI translation units.
#pragma once
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "../../include/conan_b200.h"

namespace conan {

// Thread-local message behind conan_last_error().
char* error_buffer();

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(error_buffer(), 512, fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace conan

#define CONAN_CUDA_TRY(expr)                                                              \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess)                                                            \
            return conan::fail(CONAN_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,            \
                               cudaGetErrorString(_e), __FILE__, __LINE__);               \
    } while (0)

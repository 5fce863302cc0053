// sdrx_b200 -- host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <string>

#include "../../include/sdrx_b200.h"

namespace sdrx {

std::string &last_error_ref();   // thread-local text behind sdrx_last_error()

inline int fail(int code, const std::string &msg) { last_error_ref() = msg; return code; }

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return ::sdrx::fail(SDRX_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__)); \
    } while (0)

struct DevGuard {
    int prev = -1;
    bool ok = true;
    explicit DevGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; prev = -1; cudaGetLastError(); return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) { ok = false; cudaGetLastError(); }
    }
    ~DevGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace sdrx

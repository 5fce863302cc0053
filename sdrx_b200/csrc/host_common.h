// sdrx_b200 -- host-side helpers shared by the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <map>
#include <mutex>
#include <string>
#include <utility>

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

// cudaFuncAttributeMaxDynamicSharedMemorySize is one value per function per device, shared by every engine of the
// process: only ever raise it (a second engine with a smaller N or C must not lower the limit under an earlier one's
// launches).  Call with the device current.
inline int raise_dyn_smem(const void *func, size_t bytes) {
    static std::mutex mu;
    static std::map<std::pair<int, const void *>, size_t> seen;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    size_t &cur = seen[{dev, func}];
    if (bytes <= cur) return SDRX_OK;
    CU(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    cur = bytes;
    return SDRX_OK;
}

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

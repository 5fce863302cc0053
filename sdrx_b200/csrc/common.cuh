// sdrx_b200 -- shared device helpers.
//
// "Exact" arithmetic: the generic (any-configuration) kernels restate the reference's portable
// C++ operation for operation with every multiply and add rounded on its own
// (__fmul_rn/__fadd_rn are never contracted into FMAs), so their results are bit-identical to the
// reference built without -march=native.  The specialised channel-bank kernel
// (bank_fast.cuh) trades that for FMA throughput and is held to the 1e-4 relative-RMS bar instead.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/sdrx_b200.h"

namespace sdrx {

constexpr int kChunk = SDRX_CH_IQ_BUF_SIZE;   // 512: samples per channel per block == FFT size

// ---- wire format -> iqsample_t -------------------------------------------------------------
// U8 : rtl_dev.cpp:320-323   ((float)u8) / 127.5f - 1.0f
// S16: libairspy airspy.c:58-62,:304  float = s16 / 32768 (exact: power of two)
// F32: airspy_dev.cpp:355-357 (copy)
template <int FMT> struct RawTraits;
template <> struct RawTraits<SDRX_FMT_U8>  { using vec = uchar2; static constexpr int bytes = 2; };
template <> struct RawTraits<SDRX_FMT_S16> { using vec = short2; static constexpr int bytes = 4; };
template <> struct RawTraits<SDRX_FMT_F32> { using vec = float2; static constexpr int bytes = 8; };

__host__ __device__ inline int raw_bytes_per_sample(int fmt) {
    return fmt == SDRX_FMT_U8 ? 2 : (fmt == SDRX_FMT_S16 ? 4 : 8);
}

template <int FMT>
__device__ __forceinline__ float2 convert_iq(typename RawTraits<FMT>::vec v) {
    if constexpr (FMT == SDRX_FMT_U8) {
        return make_float2(__fsub_rn(__fdiv_rn((float)v.x, 127.5f), 1.0f),
                           __fsub_rn(__fdiv_rn((float)v.y, 127.5f), 1.0f));
    } else if constexpr (FMT == SDRX_FMT_S16) {
        return make_float2((float)v.x * (1.0f / 32768.0f), (float)v.y * (1.0f / 32768.0f));
    } else {
        return v;
    }
}

template <int FMT>
__device__ __forceinline__ float2 load_iq(const void *raw, long idx) {
    using V = typename RawTraits<FMT>::vec;
    return convert_iq<FMT>(__ldg(reinterpret_cast<const V *>(raw) + idx));
}

// ---- exact complex helpers (std::complex<float> as gcc evaluates it, no contraction) --------
__device__ __forceinline__ float2 cmul_exact(float2 a, float2 b) {   // msd.hpp:134
    return make_float2(__fsub_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)),
                       __fadd_rn(__fmul_rn(a.x, b.y), __fmul_rn(a.y, b.x)));
}
__device__ __forceinline__ float norm_exact(float2 a) {               // std::norm
    return __fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y));
}
// std::abs(std::complex<float>) == glibc hypotf == (float)sqrt((double)x*x + (double)y*y)
__device__ __forceinline__ float abs_exact(float2 a) {
    double x = (double)a.x, y = (double)a.y;
    return __double2float_rn(__dsqrt_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y))));
}

__host__ __device__ inline int div_up(long a, long b) { return (int)((a + b - 1) / b); }

}  // namespace sdrx

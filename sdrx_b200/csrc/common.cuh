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

template <int FMT>
__device__ __forceinline__ float2 load_iq_from(const void *raw, long idx, int valid_lo) {
    return idx >= valid_lo ? load_iq<FMT>(raw, idx) : make_float2(0.0f, 0.0f);
}

// ---- packed FP32 pairs (sm_100a fma.rn.f32x2 family): (re, im) in one 64-bit register ----------------
typedef unsigned long long u64;
__device__ __forceinline__ u64 pack2(float x, float y) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(x), "f"(y)); return d; }
__device__ __forceinline__ float2 unpack2(u64 v) { float2 d; asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(v)); return d; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// ---- TMA bulk copies (cp.async.bulk, the 1-D form: no tensor map) completing on an mbarrier -----------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {          // make the initialised barriers visible to the async proxy
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {   // one arrival + `bytes` still to land
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {                      // one plain arrival (a consumer frees a buffer)
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    unsigned ok, spins = 0;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (!ok && ++spins > (1u << 24)) __trap();            // a copy that never lands is a bug: fail the launch, do not hang the GPU
    } while (!ok);
}
// `bytes` (a multiple of 16) from global to shared memory, both 16-byte aligned; completes `bytes` on `bar`
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// One box of a 2-D tensor map (cuTensorMapEncodeTiled) from global to shared memory; x = innermost coordinate (elements)
__device__ __forceinline__ void tma_load_2d(void *dst, const void *tmap, int x, int y, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
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

// ---- log10f exactly as the host libm computes it -------------------------------------------------
// The squelch SNR (sdrx.cpp:566) goes through glibc's log10f, which is not correctly rounded, so a
// generic device log10f differs from it in the last bit for ~5 % of arguments.  This restates glibc
// 2.39's algorithm: sysdeps/ieee754/flt-32/e_log10f.c (fdlibm split y*log10_2lo + ivln10*logf(x) +
// y*log10_2hi in float) over the table-driven logf of sysdeps/ieee754/flt-32/e_logf.c (16-entry
// table, degree-3 polynomial, double arithmetic).  tests/test_log10f.py checks the same restatement
// in C against the host's log10f over every positive float.
__device__ const double kLogfTab[16][2] = {
    {0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2}, {0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2},
    {0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2},  {0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3},
    {0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3}, {0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3},
    {0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4}, {0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4},
    {0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5}, {0x1p+0, 0x0p+0},
    {0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5},  {0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4},
    {0x1.b2036576afce6p-1, 0x1.526e57720db08p-3},  {0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3},
    {0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2},  {0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2}};

__device__ __forceinline__ float logf_glibc_normal(float x) {     // x finite, normal, > 0
    const uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.0f;
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int)tmp >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = kLogfTab[i][0], logc = kLogfTab[i][1];
    const double z = (double)__uint_as_float(iz);
    const double r = __dadd_rn(__dmul_rn(z, invc), -1.0);
    const double y0 = __dadd_rn(logc, __dmul_rn((double)k, 0x1.62e42fefa39efp-1));
    const double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(0x1.5575b0be00b6ap-2, r), -0x1.ffffef20a4123p-2);
    y = __dadd_rn(__dmul_rn(-0x1.00ea348b88334p-2, r2), y);
    y = __dadd_rn(__dmul_rn(y, r2), __dadd_rn(y0, r));
    return __double2float_rn(y);
}

__device__ __forceinline__ float log10f_glibc(float x) {
    int32_t hx = (int32_t)__float_as_uint(x);
    int32_t k = 0;
    if (hx < 0x00800000) {
        if ((hx & 0x7fffffff) == 0) return __fdiv_rn(-3.3554432000e+07f, fabsf(x));   // log(+-0) = -inf
        if (hx < 0) return __fdiv_rn(__fsub_rn(x, x), __fsub_rn(x, x));               // log(-#) = NaN
        k -= 25;
        x = __fmul_rn(x, 3.3554432000e+07f);
        hx = (int32_t)__float_as_uint(x);
    }
    if (hx >= 0x7f800000) return __fadd_rn(x, x);
    k += (hx >> 23) - 127;
    const int32_t i = (int32_t)(((uint32_t)k & 0x80000000u) >> 31);
    hx = (hx & 0x007fffff) | ((0x7f - i) << 23);
    const float y = (float)(k + i);
    const float lx = logf_glibc_normal(__uint_as_float((uint32_t)hx));
    const float z = __fadd_rn(__fmul_rn(y, 7.9034151668e-07f), __fmul_rn(4.3429449201e-01f, lx));
    return __fadd_rn(z, __fmul_rn(y, 3.0102920532e-01f));
}

// Tail fix-ups of the specialised channel bank (bank_fast.cuh): stage-3 output u of a 16-channel group
// gets up to four partial sums added, in list order.
struct TailFix { int group, u, n, src[4]; };

__host__ __device__ inline int div_up(long a, long b) { return (int)((a + b - 1) / b); }

}  // namespace sdrx

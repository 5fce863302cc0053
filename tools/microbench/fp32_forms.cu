// Instruction-form micro-benchmarks (B200 sm_100a): which FFMA/FFMA2 operand forms run at full
// rate, and how many LDS can ride along.  Feeds the design of the channel-bank kernel.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)
typedef unsigned long long u64;
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 pk(float x, float y) { u64 d; asm("mov.b64 %0, {%1,%2};" : "=l"(d) : "f"(x), "f"(y)); return d; }
__device__ __forceinline__ float2 upk(u64 v) { float2 d; asm("mov.b64 {%0,%1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(v)); return d; }

// (1) 3 distinct register sources, no reuse possible: acc[i] = fma(a[i], b[i], acc[i])
__global__ void __launch_bounds__(512) k_ffma_3reg(float *out, int iters) {
    float acc[8], a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] = 0.f; a[i] = threadIdx.x * 1e-3f + i; b[i] = 1.f - i * 1e-3f; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(acc[i]) : "f"(a[i]), "f"(b[i]));
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// (2) FFMA2, 3 distinct register-pair sources
__global__ void __launch_bounds__(512) k_ffma2_3reg(float *out, int iters) {
    u64 acc[8], a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] = 0; a[i] = pk(threadIdx.x * 1e-3f + i, 0.5f + i); b[i] = pk(1.f - i * 1e-3f, 0.25f); }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] = fma2(a[i], b[i], acc[i]);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 v = upk(acc[i]); s += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// (3) the mix + FIR shape: per "sample": 1 FMUL2 + 1 FFMA2 (3 reg) then NT FFMA2 (reg, reg(hh), acc)
template <int NT>
__global__ void __launch_bounds__(512) k_mixfir2(float *out, int iters, const float *prm) {
    u64 acc[4] = {0, 0, 0, 0};
    u64 xa = pk(threadIdx.x * 1e-3f, threadIdx.x * 1e-3f), xb = pk(prm[0], prm[0]);
    u64 ta = pk(prm[1], prm[2]), tb = pk(-prm[2], prm[1]);
    u64 hh[4] = {pk(prm[3], prm[3]), pk(prm[4], prm[4]), pk(prm[5], prm[5]), pk(prm[6], prm[6])};
    for (int it = 0; it < iters; ++it) {
        u64 s = mul2(xa, ta);
        s = fma2(xb, tb, s);
#pragma unroll
        for (int i = 0; i < NT; ++i) acc[i & 3] = fma2(s, hh[i & 3], acc[i & 3]);
        xa = acc[0];
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 v = upk(acc[i]); sum += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}
// (4) scalar mix (4 x 3-reg ops) + NT FFMA2
template <int NT>
__global__ void __launch_bounds__(512) k_mixs_fir2(float *out, int iters, const float *prm) {
    u64 acc[4] = {0, 0, 0, 0};
    float xr = threadIdx.x * 1e-3f, xi = prm[0], tr = prm[1], ti = prm[2];
    u64 hh[4] = {pk(prm[3], prm[3]), pk(prm[4], prm[4]), pk(prm[5], prm[5]), pk(prm[6], prm[6])};
    for (int it = 0; it < iters; ++it) {
        float sr, si;
        asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(sr) : "f"(xi), "f"(ti));
        asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(sr) : "f"(xr), "f"(tr), "f"(-sr));
        asm volatile("mul.rn.f32 %0, %1, %2;" : "=f"(si) : "f"(xi), "f"(tr));
        asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(si) : "f"(xr), "f"(ti), "f"(si));
        u64 s = pk(sr, si);
#pragma unroll
        for (int i = 0; i < NT; ++i) acc[i & 3] = fma2(s, hh[i & 3], acc[i & 3]);
        xr = upk(acc[0]).x;
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 v = upk(acc[i]); sum += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
}
// (5) FFMA2 stream with K LDS.128 per 16 FFMA2 (smem co-issue headroom when FMA-pipe-bound)
template <int NL>
__global__ void __launch_bounds__(512) k_ffma2_lds(float *out, int iters) {
    __shared__ float4 sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_float4(i * 1e-4f, 0.5f, 0.25f, 1.f);
    __syncthreads();
    u64 acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = pk(i, i + 1);
    u64 hh = pk(0.999f, 0.999f);
    int base = (threadIdx.x & 7);    // 8 distinct 16-byte addresses per warp -> 1 wavefront
    for (int it = 0; it < iters; ++it) {
        float4 v[NL > 0 ? NL : 1];
#pragma unroll
        for (int l = 0; l < NL; ++l) v[l] = sm[(base + it * 8 + l * 64) & 2047];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            u64 s = (NL > 0) ? ((i & 1) ? pk(v[i % (NL > 0 ? NL : 1)].x, v[i % (NL > 0 ? NL : 1)].y) : pk(v[i % (NL > 0 ? NL : 1)].z, v[i % (NL > 0 ? NL : 1)].w)) : hh;
            acc[i & 7] = fma2(acc[i & 7], hh, s);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float2 v = upk(acc[i]); s += v.x + v.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F> float time_ms(F f) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) { CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms; }
    return best;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount; const int iters = 20000, threads = 512; int blocks = sms * 4;
    float *out; CK(cudaMalloc(&out, sizeof(float) * blocks * threads));
    double lanes = (double)blocks * threads;
    float hp[8] = {0.3f, 0.9f, 0.1f, 0.1f, 0.2f, 0.3f, 0.4f, 0.f}; float *prm; CK(cudaMalloc(&prm, sizeof(hp))); CK(cudaMemcpy(prm, hp, sizeof(hp), cudaMemcpyHostToDevice));
    auto rep = [&](const char *name, float ms, double flops_per_iter, double instr_per_iter) {
        printf("%-28s %.3f ms  %.2f TFLOP/s  %.2f Tinstr-lane/s\n", name, ms, lanes * iters * flops_per_iter / ms * 1e-9, lanes * iters * instr_per_iter / ms * 1e-9);
    };
    rep("ffma 3reg", time_ms([&] { k_ffma_3reg<<<blocks, threads>>>(out, iters); }), 16, 8);
    rep("ffma2 3reg", time_ms([&] { k_ffma2_3reg<<<blocks, threads>>>(out, iters); }), 32, 8);
    rep("mix2+fir2 NT=3.4~4", time_ms([&] { k_mixfir2<4><<<blocks, threads>>>(out, iters, prm); }), 6 * 4, 6);
    rep("mix2+fir2 NT=8", time_ms([&] { k_mixfir2<8><<<blocks, threads>>>(out, iters, prm); }), 10 * 4, 10);
    rep("mixs+fir2 NT=4", time_ms([&] { k_mixs_fir2<4><<<blocks, threads>>>(out, iters, prm); }), 4 * 2 + 4 * 4, 8);
    rep("mixs+fir2 NT=8", time_ms([&] { k_mixs_fir2<8><<<blocks, threads>>>(out, iters, prm); }), 4 * 2 + 8 * 4, 12);
    rep("ffma2x16 + 0 lds", time_ms([&] { k_ffma2_lds<0><<<blocks, threads>>>(out, iters); }), 64, 16);
    rep("ffma2x16 + 2 lds.128", time_ms([&] { k_ffma2_lds<2><<<blocks, threads>>>(out, iters); }), 64, 18);
    rep("ffma2x16 + 4 lds.128", time_ms([&] { k_ffma2_lds<4><<<blocks, threads>>>(out, iters); }), 64, 20);
    rep("ffma2x16 + 8 lds.128", time_ms([&] { k_ffma2_lds<8><<<blocks, threads>>>(out, iters); }), 64, 24);
    return 0;
}

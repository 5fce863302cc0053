// FP32 CUDA-core peak micro-benchmark for B200 (sm_100a): the roofline denominator for the
// channel-bank kernel ("measure it with an FMA micro-benchmark on the box").
// Variants: scalar FFMA (3-reg), FFMA with constant-bank operand, packed fma.rn.f32x2,
// and FFMA interleaved with shared-memory loads (issue-slot pressure).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__constant__ float c_taps[32];

template <int ILP>
__global__ void __launch_bounds__(1024) k_ffma(float *out, int iters, float a, float b) {
    float acc[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) acc[i] = threadIdx.x * 0.001f + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// acc += x * c[k]  (constant-bank operand, distinct x registers): the FIR inner-loop shape
template <int ILP>
__global__ void __launch_bounds__(1024) k_ffma_const(float *out, int iters, float a) {
    float acc[ILP], x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) { acc[i] = 0.f; x[i] = threadIdx.x * 0.001f + i + a; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = fmaf(x[(i + 1) % ILP], c_taps[i], acc[i]);
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = fmaf(x[(i + 3) % ILP], c_taps[i + 8], acc[i]);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d;
    asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int ILP>
__global__ void __launch_bounds__(1024) k_ffma2(float *out, int iters, float a, float b) {
    unsigned long long acc[ILP];
    unsigned long long aa, bb;
    asm("mov.b64 %0, {%1, %2};" : "=l"(aa) : "f"(a), "f"(a * 1.0001f));
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(b), "f"(b * 1.0001f));
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        float v = threadIdx.x * 0.001f + i;
        asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(v), "f"(v + 0.5f));
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) acc[i] = fma2(acc[i], aa, bb);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
        s += lo + hi;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// FFMA : LDS mix.  Per iteration: NF ffma + 1 LDS.64 (lane-distinct, conflict free).
template <int NF>
__global__ void __launch_bounds__(1024) k_ffma_lds(float *out, int iters, float a, float b) {
    __shared__ float2 sm[2048];
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) sm[i] = make_float2(i * 1e-3f, i * 2e-3f);
    __syncthreads();
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = threadIdx.x * 0.001f + i;
    int idx = threadIdx.x;
    for (int it = 0; it < iters; ++it) {
        float2 v = sm[(idx + it * 32) & 2047];
#pragma unroll
        for (int i = 0; i < NF; ++i) acc[i % 8] = fmaf(acc[i % 8], (i & 1) ? v.x : v.y, b);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
float time_ms(F f, int reps = 5) {
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f(); f();
    CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < reps; ++r) {
        CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
    }
    return best;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount;
    printf("device %s sms %d clock %d kHz\n", p.name, sms, p.clockRate);
    float h[32]; for (int i = 0; i < 32; ++i) h[i] = 0.01f * (i + 1);
    CK(cudaMemcpyToSymbol(c_taps, h, sizeof(h)));
    float *out; CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 1024));
    const int iters = 20000;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        double lanes = (double)blocks * threads;
        { float ms = time_ms([&] { k_ffma<8><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma      ILP8  thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 8 * 2 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma<16><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma      ILP16 thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 16 * 2 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma_const<8><<<blocks, threads>>>(out, iters, 0.5f); });
          printf("ffma_c[]  ILP8  thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 16 * 2 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma2<8><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma2     ILP8  thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 8 * 4 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma2<16><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma2     ILP16 thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 16 * 4 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma_lds<8><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma8+lds       thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 8 * 2 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma_lds<4><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma4+lds       thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 4 * 2 / ms * 1e-9); }
        { float ms = time_ms([&] { k_ffma_lds<16><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f); });
          printf("ffma16+lds      thr %4d: %.3f ms  %.2f TFLOP/s\n", threads, ms, lanes * iters * 16 * 2 / ms * 1e-9); }
    }
    // sustained: 3 s of back-to-back FFMA launches
    {
        int threads = 512, blocks = sms * 4;
        cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0));
        int n = 0; float ms = 0;
        for (; n < 4000; ++n) {
            k_ffma<8><<<blocks, threads>>>(out, iters, 1.0001f, 0.5f);
            if ((n & 31) == 31) { CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms > 3000) { ++n; break; } }
        }
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("ffma sustained %d launches %.1f ms: %.2f TFLOP/s\n", n, ms, (double)blocks * threads * iters * 8 * 2 * n / ms * 1e-9);
    }
    CK(cudaFree(out));
    return 0;
}

// sdrx_b200 -- polyphase / DFT channelizer for the uniform 8.33 kHz grid (the direction the
// reference's README names: "not the most efficient way when listening to many simultaneous channels", README.md:25-28).
//
// OPT-IN, NOT THE PARITY PATH (SDRX_FLAG_POLYPHASE).  In exact arithmetic it computes the same map as the reference chain:
// with h_c = h1(z) h2(z^5) h3(z^25) h4(z^125) (the four MSD stages of msd.hpp:251-335 as one 8857-tap response, decimation
// D = 625) and W = exp(-2 pi i / N), channel k's output M is
//     y_k[M] = sum_j h_c[j] x[n_M - j] W^(k (n_M - j)),   n_M = D M + D - 1 (absolute sample index),
// and because W^(k n) only depends on n mod N (N = 1200: the translator period, sdrx.cpp:1571-1583), folding the windowed
// input into N bins first, u[p] = sum over {n = p mod N} h_c[n_M - n] x[n], turns ALL channels' outputs at instant M into one
// N-point DFT: y_k[M] = DFT_N(u)[k mod N].  17 k complex MACs per output instant for every channel together instead of
// 8857 per channel.  It is not bit-compatible: the reference mixes with cexpf of a float-ROUNDED angle (sdrx.cpp:1682-1690,
// up to 1.2e-4 rad off at |k| = 480), a per-sample phase jitter on the whole band that exact twiddles do not reproduce;
// tests/test_gpu_polyphase.py and bench.py --polyphase measure what that costs per channel.
//
// One CTA per output instant: fold (coalesced reads of the 8857-sample window, 7-8 taps per bin), then the 1200-point DFT
// as 40 x 30 Cooley-Tukey in shared memory (step A: 30 DFTs of 40 points, twiddle, step B: 30-point DFTs evaluated only at
// the bins that are listened to), FP32 FMA, twiddles from one table of exp(-2 pi i n / 1200) rounded once from double.
#pragma once
#include "common.cuh"

namespace sdrx {

constexpr int kPolyN = 1200;          // DFT size == translator period at 10 MS/s
constexpr int kPolyN1 = 40, kPolyN2 = 30;
constexpr int kPolyThreads = 256;
inline size_t poly_smem_bytes() { return sizeof(float2) * 3 * kPolyN; }

template <int FMT>
static __global__ void __launch_bounds__(kPolyThreads)
k_poly_channelizer(const void *__restrict__ raw, int hist, int valid_lo, int pos0 /* trans_pos at the first sample of the block */,
                   int D, const float *__restrict__ hc, int Lh, const float2 *__restrict__ wtab, const int *__restrict__ bin, int C,
                   float2 *__restrict__ ch_iq, long iq_stride) {
    extern __shared__ __align__(16) float2 psm[];
    float2 *w = psm;                  // [1200] exp(-2 pi i n / 1200)
    float2 *u = psm + kPolyN;         // [1200] folded window
    float2 *a = psm + 2 * kPolyN;     // [30][40] after step A and the twiddle
    const int t = threadIdx.x;
    const int M = blockIdx.x;
    for (int i = t; i < kPolyN; i += kPolyThreads) w[i] = wtab[i];
    // ---- fold: bin p collects the samples whose absolute index is p mod N, newest first
    const int nM = D * M + D - 1;                          // block-relative index of the sample that completes output M
    for (int p = t; p < kPolyN; p += kPolyThreads) {
        int d = (pos0 + nM - p) % kPolyN;
        if (d < 0) d += kPolyN;                            // distance from n_M back to the newest sample of bin p
        float2 acc = make_float2(0.f, 0.f);
        for (int j = d; j < Lh; j += kPolyN) {
            const float2 x = load_iq_from<FMT>(raw, (long)hist + nM - j, valid_lo);
            const float h = hc[j];
            acc.x = fmaf(h, x.x, acc.x);
            acc.y = fmaf(h, x.y, acc.y);
        }
        u[p] = acc;
    }
    __syncthreads();
    // ---- step A: A[n2][k1] = W1200^(n2 k1) * sum_n1 W40^(n1 k1) u[30 n1 + n2]      (W40^(n1 k1) = W1200^(30 n1 k1))
    for (int idx = t; idx < kPolyN; idx += kPolyThreads) {
        const int n2 = idx / kPolyN1, k1 = idx - n2 * kPolyN1;
        float2 acc = make_float2(0.f, 0.f);
        int wi = 0;                                        // (30 n1 k1) mod 1200
        const int wstep = (kPolyN2 * k1) % kPolyN;
#pragma unroll 8
        for (int n1 = 0; n1 < kPolyN1; ++n1) {
            const float2 x = u[kPolyN2 * n1 + n2], ww = w[wi];
            acc.x = fmaf(x.x, ww.x, fmaf(-x.y, ww.y, acc.x));
            acc.y = fmaf(x.x, ww.y, fmaf(x.y, ww.x, acc.y));
            wi += wstep;
            if (wi >= kPolyN) wi -= kPolyN;
        }
        const float2 tw = w[(n2 * k1) % kPolyN];
        a[idx] = make_float2(acc.x * tw.x - acc.y * tw.y, acc.x * tw.y + acc.y * tw.x);
    }
    __syncthreads();
    // ---- step B, only at the bins that are listened to: X[k1 + 40 k2] = sum_n2 W30^(n2 k2) A[n2][k1]   (W30 = W1200^40)
    for (int c = t; c < C; c += kPolyThreads) {
        const int k = bin[c];
        const int k2 = k / kPolyN1, k1 = k - k2 * kPolyN1;
        float2 acc = make_float2(0.f, 0.f);
        int wi = 0;
        const int wstep = (kPolyN1 * k2) % kPolyN;
#pragma unroll 6
        for (int n2 = 0; n2 < kPolyN2; ++n2) {
            const float2 x = a[n2 * kPolyN1 + k1], ww = w[wi];
            acc.x = fmaf(x.x, ww.x, fmaf(-x.y, ww.y, acc.x));
            acc.y = fmaf(x.x, ww.y, fmaf(x.y, ww.x, acc.y));
            wi += wstep;
            if (wi >= kPolyN) wi -= kPolyN;
        }
        ch_iq[(long)c * iq_stride + M] = acc;
    }
}

}  // namespace sdrx

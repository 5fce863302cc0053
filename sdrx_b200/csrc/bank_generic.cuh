// sdrx_b200 -- generic channelizer kernels: any stage list, any block length, exact arithmetic.
//
// These restate MSD::decimate (msd.hpp:75-159) stage by stage.  Instead of the reference's
// per-sample delay-line walk, stage output k is computed directly from its input window
// (SURVEY.md §3.4): it is emitted after the stage's input index first + k*m and equals
//   sum_{t<H} h[t] * x[first + k*m - (H-1) + t]
// with the stage's last H-1 inputs of the previous call kept as a history prefix of the input
// buffer (zeros before the stream start, msd.hpp:167).  The summation itself follows the
// reference's portable folded loop (msd.hpp:308-331) add for add, so results are bit-identical.
#pragma once

#include "common.cuh"

namespace sdrx {

// Stage 0: raw -> float, NCO mix (msd.hpp:134-135) and the first decimating FIR, all channels.
//   raw          [hist + n_in] samples in wire format; raw[hist + i] is call-relative input i
//   table        [N][Cpad] translator, transposed so that a warp of 32 channels reads 256 B
//   has_tr[c]    0 -> the reference's no-translator path (msd.hpp:79-101), no multiply at all
//   pos0         trans_pos_ of call-relative input 0
//   valid_lo     buffer index of the first sample of the stream still in the history prefix; anything
//                before it is the reference's zero-initialised delay line (msd.hpp:167) -- float 0, which
//                a zero *byte* is not in the u8 wire format
// grid (ceil(n_out/8), Cpad/32), block (32, 8): lane = channel, threadIdx.y = output.
template <int FMT>
static __global__ void __launch_bounds__(256)
k_mix_fir_generic(const void *__restrict__ raw, int hist, const float2 *__restrict__ table, int Cpad, int C,
                  int N, int pos0, const uint8_t *__restrict__ has_tr, const float *__restrict__ h, int H, int m,
                  int first, int n_out, float2 *__restrict__ out, long out_stride, int out_off, int valid_lo) {
    const int c = blockIdx.y * 32 + threadIdx.x;
    const int k = blockIdx.x * 8 + threadIdx.y;
    if (k >= n_out || c >= C) return;
    const int half = (H - 1) >> 1;
    const long i0 = (long)first + (long)k * m - (H - 1);     // call-relative index of window tap 0
    const bool tr_on = has_tr[c] != 0;
    int idx_lo = (int)(((long)pos0 + i0) % N);
    if (idx_lo < 0) idx_lo += N;
    int idx_hi = (int)(((long)pos0 + i0 + (H - 1)) % N);
    if (idx_hi < 0) idx_hi += N;
    const long p_lo = hist + i0;

    float re = 0.0f, im = 0.0f;
    for (int i = 0; i < half; ++i) {
        float2 a = load_iq_from<FMT>(raw, p_lo + i, valid_lo);
        float2 b = load_iq_from<FMT>(raw, p_lo + (H - 1) - i, valid_lo);
        if (tr_on) {
            a = cmul_exact(a, __ldg(table + (long)idx_lo * Cpad + c));
            b = cmul_exact(b, __ldg(table + (long)idx_hi * Cpad + c));
        }
        const float hv = __ldg(h + i);
        re = __fadd_rn(re, __fmul_rn(__fadd_rn(a.x, b.x), hv));
        im = __fadd_rn(im, __fmul_rn(__fadd_rn(a.y, b.y), hv));
        if (++idx_lo == N) idx_lo = 0;
        if (--idx_hi < 0) idx_hi = N - 1;
    }
    {
        float2 a = load_iq_from<FMT>(raw, p_lo + half, valid_lo);
        if (tr_on) a = cmul_exact(a, __ldg(table + (long)idx_lo * Cpad + c));
        const float hv = __ldg(h + half);
        re = __fadd_rn(re, __fmul_rn(a.x, hv));
        im = __fadd_rn(im, __fmul_rn(a.y, hv));
    }
    out[(long)c * out_stride + out_off + k] = make_float2(re, im);
}

// Stage 0 of the --ftfir path (msd.hpp:104-131): the raw samples enter the first stage's delay line unmixed and
// its output is S::calculateOutputTranslated (msd.hpp:341-423, portable loop :398-416): a complex x complex dot
// product with the coefficient set hk_[k_], k_ advancing by one per output and wrapping at N/m sets (:420).
//   hk      [n_sets * H][Cpad] float2, hk[(set * H + i) * Cpad + c] = h[i] * tr_c[(j0 + set*m + i) % N] * 2  (msd.hpp:192-224)
//   k0      k_ before the call's first output
// Channels without a translator take the plain folded stage (msd.hpp:79-101), as in the reference.
// grid (ceil(n_out/8), Cpad/32), block (32, 8): lane = channel, threadIdx.y = output.
template <int FMT>
static __global__ void __launch_bounds__(256)
k_ftfir_generic(const void *__restrict__ raw, int hist, const float2 *__restrict__ hk, int n_sets, int k0, int Cpad, int C,
                const uint8_t *__restrict__ has_tr, const float *__restrict__ h, int H, int m, int first, int n_out,
                float2 *__restrict__ out, long out_stride, int out_off, int valid_lo) {
    const int c = blockIdx.y * 32 + threadIdx.x;
    const int k = blockIdx.x * 8 + threadIdx.y;
    if (k >= n_out || c >= C) return;
    const long p_lo = hist + (long)first + (long)k * m - (H - 1);
    float re = 0.0f, im = 0.0f;
    if (has_tr[c]) {
        const float2 *hs = hk + ((long)((k0 + k) % n_sets) * H) * Cpad + c;
        for (int i = 0; i < H; ++i) {
            const float2 d = load_iq_from<FMT>(raw, p_lo + i, valid_lo);
            const float2 g = __ldg(hs + (long)i * Cpad);
            re = __fadd_rn(re, __fsub_rn(__fmul_rn(d.x, g.x), __fmul_rn(d.y, g.y)));
            im = __fadd_rn(im, __fadd_rn(__fmul_rn(d.x, g.y), __fmul_rn(d.y, g.x)));
        }
    } else {
        const int half = (H - 1) >> 1;
        for (int i = 0; i < half; ++i) {
            const float2 a = load_iq_from<FMT>(raw, p_lo + i, valid_lo);
            const float2 b = load_iq_from<FMT>(raw, p_lo + (H - 1) - i, valid_lo);
            const float hv = __ldg(h + i);
            re = __fadd_rn(re, __fmul_rn(__fadd_rn(a.x, b.x), hv));
            im = __fadd_rn(im, __fmul_rn(__fadd_rn(a.y, b.y), hv));
        }
        const float2 a = load_iq_from<FMT>(raw, p_lo + half, valid_lo);
        const float hv = __ldg(h + half);
        re = __fadd_rn(re, __fmul_rn(a.x, hv));
        im = __fadd_rn(im, __fmul_rn(a.y, hv));
    }
    out[(long)c * out_stride + out_off + k] = make_float2(re, im);
}

// Stage s >= 1: decimating symmetric FIR on a per-channel stream.
//   in  [C][in_stride], in[c][in_hist + i] is call-relative input i (in_hist == H-1)
// grid (ceil(n_out/128), C), block 128: lane = output.
static __global__ void __launch_bounds__(128)
k_fir_generic(const float2 *__restrict__ in, long in_stride, int in_hist, const float *__restrict__ h, int H, int m,
              int first, int n_out, float2 *__restrict__ out, long out_stride, int out_off) {
    const int c = blockIdx.y;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_out) return;
    const int half = (H - 1) >> 1;
    const float2 *d = in + (long)c * in_stride + in_hist + (long)first + (long)k * m - (H - 1);
    float re = 0.0f, im = 0.0f;
    for (int i = 0; i < half; ++i) {
        const float2 a = d[i], b = d[H - 1 - i];
        const float hv = __ldg(h + i);
        re = __fadd_rn(re, __fmul_rn(__fadd_rn(a.x, b.x), hv));
        im = __fadd_rn(im, __fmul_rn(__fadd_rn(a.y, b.y), hv));
    }
    const float2 a = d[half];
    const float hv = __ldg(h + half);
    re = __fadd_rn(re, __fmul_rn(a.x, hv));
    im = __fadd_rn(im, __fmul_rn(a.y, hv));
    out[(long)c * out_stride + out_off + k] = make_float2(re, im);
}

// Carry the last `hist` entries of [hist | n_in new] to the front of the same per-channel buffer
// (the reference's delay line, msd.hpp:228-247, seen as a sliding prefix).  One CTA per channel;
// reads complete before any write, so n_in < hist (overlap) is safe.  hist <= 1024.
static __global__ void __launch_bounds__(256)
k_carry_hist(float2 *__restrict__ buf, long stride, int hist, int n_in) {
    float2 *b = buf + (long)blockIdx.x * stride;
    float2 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int i = threadIdx.x + j * 256;
        if (i < hist) v[j] = b[n_in + i];
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int i = threadIdx.x + j * 256;
        if (i < hist) b[i] = v[j];
    }
}

// Block preparation, one launch:
//   CTAs [0, n_power): block power pass 1 -- deterministic per-CTA partial sums of re^2+im^2
//                      (rtl_dev.cpp:333-337, airspy_dev.cpp:366-405); the mean / 10*log10f - 3 is taken
//                      by the mix kernel from the partials in a fixed order.
//   CTAs [n_power, ..): the raw delay line between the two ping-pong input buffers:
//                      dst[0..hist) = src[n_in .. n_in+hist), byte granular (src holds [hist | n_in]).
// The power CTAs only read dst[hist..], the carry CTAs only write dst[0..hist): no hazard.
// blockIdx.y = stream of a multi-stream engine (each stream's [hist | block] sits stream_bytes further; partials n_power apart).
template <int FMT>
static __global__ void __launch_bounds__(256)
k_prep(const void *__restrict__ raw_, int hist, int n, float *__restrict__ partial_, int n_power,
       uint8_t *__restrict__ dst_, const uint8_t *__restrict__ src_, long n_in_bytes, int hist_bytes, long stream_bytes) {
    const void *raw = reinterpret_cast<const uint8_t *>(raw_) + blockIdx.y * stream_bytes;
    uint8_t *dst = dst_ + blockIdx.y * stream_bytes;
    const uint8_t *src = src_ + blockIdx.y * stream_bytes;
    float *partial = partial_ + blockIdx.y * n_power;
    if ((int)blockIdx.x >= n_power) {
        const int i = ((int)blockIdx.x - n_power) * 256 + threadIdx.x;
        if (i < hist_bytes) dst[i] = src[n_in_bytes + i];
        return;
    }
    __shared__ float sh[256];
    float acc = 0.0f;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += n_power * 256) {
        const float2 v = load_iq<FMT>(raw, (long)hist + i);
        acc += v.x * v.x + v.y * v.y;
    }
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

}  // namespace sdrx

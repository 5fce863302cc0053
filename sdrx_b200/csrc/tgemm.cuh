// sdrx_b200 -- boundary A as "fold + table GEMM": the whole multistage channelizer (msd.hpp:133-154, :251-335) of every
// channel as ONE matrix product per block, with the reference's own translator tables as the matrix.
//
// The chain of a channel is linear and its translator is periodic (period N, sdrx.cpp:1571-1583), so with the stages
// collapsed into one prototype h_c(z) = h_1(z) h_2(z^m1) h_3(z^(m1 m2)) ... (8857 taps, decimation D = 625 at 10 MS/s)
//     y_c[M] = sum_j h_c[j] x[n_M - j] T_c[(n_M - j) mod N],     n_M = D M + D - 1  (absolute sample index)
//            = sum_{p < N} T_c[p] * u_M[p],      u_M[p] = sum over {n = p mod N} h_c[n_M - n] x[n]      ("the fold")
// for ANY table T_c -- no property of exp() is used, so T_c is the reference's own float table (cexpf of a float-rounded
// angle, sdrx.cpp:1682-1690) and the result differs from the reference only by the order of the float additions, like the
// time-domain bank (1e-4 bar, measured ~1e-5), not by the 3e-6-of-the-band phase jitter an exact-twiddle FFT channelizer
// leaves out (polyphase.cuh).  Y[C x outputs] = T[C x N] U[N x outputs]: a complex GEMM, which is what north_star's
// "recast as a complex GEMM (channels x taps)" asks for; u_M is shared by all channels, so the per-channel work drops from
// 19.2 flop per input sample to 8 N / D = 15.4 flop per input sample -- and to half of that for +-offset pairs, whose
// tables are complex conjugates (bit for bit, see bank_fast.cuh): with P = Tr Ur, Q = Ti Ui, R = Tr Ui, S = Ti Ur summed over p,
//     Y_{+k} = (P - Q, R + S),   Y_{-k} = (P + Q, R - S).
//
//   k_fold<FMT>   u_M[p] for kFoldM consecutive output instants per CTA: coalesced reads of the raw window (wire format,
//                 converted on the fly), 7-8 taps per bin, written as U[p][M] (float2, M contiguous).
//   k_tgemm       FP32 CUDA-core GEMM with packed FFMA2: a CTA owns a 32-slot x 64-instant tile of Y, walks p in chunks
//                 of 16 through double-buffered cp.async tiles; a thread owns 4 slots x 4 instants = 32 packed
//                 accumulators ((P,R) and (S,Q) per output) and per p issues 4 LDS.128 (table: (Tr,Tr,Ti,Ti), a broadcast),
//                 2 LDS.128 (U) and 32 FFMA2.  Epilogue: un-mix the pair and store both channels' rows.
#pragma once
#include <vector>

#include "common.cuh"
#include "host_common.h"

namespace sdrx {

constexpr int kFoldM = 2;             // output instants per fold CTA
constexpr int kFoldThreads = 256;
constexpr int kGemmSlots = 32;        // tile: pair slots
constexpr int kGemmM = 64;            //       output instants
constexpr int kGemmK = 16;            // table period entries per staged chunk == terms of one partial sum
constexpr int kGemmBufs = 3;          // staged chunks in flight
constexpr int kGemmThreads = 256;     // 8 x 32 threads of 4 slots x 2 instants
constexpr size_t kGemmSmem = kGemmBufs * (sizeof(float4) * kGemmK * kGemmSlots + sizeof(float2) * kGemmK * kGemmM);   // 48 KB

// U[stream][p][Mpad] (float2).  grid (ceil(n_out / kFoldM), n_streams).  x index of output M's newest sample: hist + D M + D - 1.
template <int FMT>
static __global__ void __launch_bounds__(kFoldThreads)
k_fold(const void *__restrict__ raw_all, long stream_bytes, int hist, int valid_lo, int pos0, int D, int N, const float *__restrict__ hc, int Lh,
       int n_out, float2 *__restrict__ U, int Mpad) {
    const void *raw = reinterpret_cast<const unsigned char *>(raw_all) + (long)blockIdx.y * stream_bytes;
    float2 *Us = U + (long)blockIdx.y * N * Mpad;
    const int M0 = blockIdx.x * kFoldM;
    // items (m, p), p fastest: consecutive threads read consecutive samples
    for (int it = threadIdx.x; it < kFoldM * N; it += kFoldThreads) {
        const int m = it / N, p = it - m * N;
        const int M = M0 + m;
        if (M >= n_out) break;
        const int nM = D * M + D - 1;
        int d = (pos0 + nM - p) % N;
        if (d < 0) d += N;                                  // distance from n_M back to the newest sample of bin p
        float2 acc = make_float2(0.f, 0.f);
        for (int j0 = d; j0 < Lh; j0 += 8 * N) {           // eight taps at a time: all their loads in flight before the sum
            float2 xv[8];
            float hv[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int j = j0 + q * N;
                const bool in = j < Lh;
                xv[q] = in ? load_iq_from<FMT>(raw, (long)hist + nM - j, valid_lo) : make_float2(0.f, 0.f);
                hv[q] = in ? __ldg(hc + j) : 0.0f;
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) {                  // newest sample first, as the loop it unrolls
                acc.x = fmaf(hv[q], xv[q].x, acc.x);
                acc.y = fmaf(hv[q], xv[q].y, acc.y);
            }
        }
        Us[(long)p * Mpad + M] = acc;
    }
}

// Y = T U for one block.  Tdup[p][slots_pad] = (Tr, Tr, Ti, Ti) of the slot's first channel; slot_chan[slot][2] = the
// channel with that table and its conjugate partner (-1: none).  grid (slots_pad / 32, Mpad / 64, n_streams).
//
// Summation order: the 1200 terms of an output are NOT added in one running sum.  A weak channel next to strong carriers
// is the small difference of large oscillating partial sums, and a single FP32 running sum loses it (measured against the
// oracle: 14 of 960 channels above the 1e-4 bar, worst 1.6e-4); partial sums over the 16 terms of a staged chunk, added
// chunk by chunk into the total, keep every partial small: worst channel 4.7e-5, the level of the time-domain bank.
static __global__ void __launch_bounds__(kGemmThreads)
k_tgemm(const float4 *__restrict__ Tdup, int slots_pad, const float2 *__restrict__ U, int Mpad, int N, const int *__restrict__ slot_chan,
        int chans_per_stream, float2 *__restrict__ ch_iq, long iq_stride, int n_out) {
    extern __shared__ __align__(16) unsigned char gsm[];
    float4 *As = reinterpret_cast<float4 *>(gsm);                                                     // [bufs][kGemmK][kGemmSlots]
    float2 *Bs = reinterpret_cast<float2 *>(gsm + kGemmBufs * sizeof(float4) * kGemmK * kGemmSlots);  // [bufs][kGemmK][kGemmM]
    const int tid = threadIdx.x;
    const int ts = tid >> 5, tm = tid & 31;               // 4 slots from ts * 4, 2 instants from tm * 2
    const int s0 = blockIdx.x * kGemmSlots, m0 = blockIdx.y * kGemmM;
    const float2 *Us = U + (long)blockIdx.z * N * Mpad;

    auto stage = [&](int chunk) {                          // 16 table rows of 32 slots (512 B each), 16 U rows of 64 instants (512 B each)
        const int p0 = chunk * kGemmK, buf = chunk % kGemmBufs;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            const int i = tid + q * kGemmThreads;          // 0..511: 16-byte pieces
            const int r = i >> 5, c = i & 31;
            const unsigned da = smem_u32(As + (buf * kGemmK + r) * kGemmSlots + c);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(da), "l"(Tdup + (long)(p0 + r) * slots_pad + s0 + c) : "memory");
            const unsigned db = smem_u32(Bs + (buf * kGemmK + r) * kGemmM + 2 * c);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(db), "l"(Us + (long)(p0 + r) * Mpad + m0 + 2 * c) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    u64 tpr[4][2], tsq[4][2];                              // totals [slot][instant]: (P, R) and (S, Q)
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 2; ++b) tpr[a][b] = tsq[a][b] = 0ull;
    const int n_chunks = N / kGemmK;
    stage(0);
    if (n_chunks > 1) stage(1); else asm volatile("cp.async.commit_group;" ::: "memory");
    for (int ch = 0; ch < n_chunks; ++ch) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");   // chunk ch has landed (chunk ch + 1 may be in flight)
        __syncthreads();                                   // ... for every thread; and every thread is done with chunk ch - 1
        if (ch + 2 < n_chunks) stage(ch + 2);              // into the buffer chunk ch - 1 was read from
        else asm volatile("cp.async.commit_group;" ::: "memory");
        const float4 *Ab = As + (ch % kGemmBufs) * kGemmK * kGemmSlots + ts * 4;
        const float4 *Bb = reinterpret_cast<const float4 *>(Bs + (ch % kGemmBufs) * kGemmK * kGemmM + tm * 2);
        u64 pr[4][2], sq[4][2];                            // this chunk's partial sums
#pragma unroll
        for (int k = 0; k < kGemmK; ++k) {
            const float4 u01 = Bb[k * (kGemmM / 2)];
            const u64 uu[2] = {pack2(u01.x, u01.y), pack2(u01.z, u01.w)};
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const float4 t4 = Ab[k * kGemmSlots + a];
                const u64 trr = pack2(t4.x, t4.y), tii = pack2(t4.z, t4.w);
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    if (k == 0) { pr[a][b] = mul2(uu[b], trr); sq[a][b] = mul2(uu[b], tii); }
                    else {
                        pr[a][b] = fma2(uu[b], trr, pr[a][b]);   // (Tr Ur, Tr Ui) = (P, R)
                        sq[a][b] = fma2(uu[b], tii, sq[a][b]);   // (Ti Ur, Ti Ui) = (S, Q)
                    }
                }
            }
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) { tpr[a][b] = add2(tpr[a][b], pr[a][b]); tsq[a][b] = add2(tsq[a][b], sq[a][b]); }
    }
    // ---- epilogue: Y_first = (P - Q, R + S), Y_partner = (P + Q, R - S); two instants = one 16-byte store per channel
    const long coff = (long)blockIdx.z * chans_per_stream;
    const int M = m0 + tm * 2;
    if (M < n_out) {
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int slot = s0 + ts * 4 + a;
            const int c1 = slot_chan[2 * slot], c2 = slot_chan[2 * slot + 1];
            float2 y1[2], y2[2];
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const float2 v = unpack2(tpr[a][b]), w = unpack2(tsq[a][b]);     // v = (P, R), w = (S, Q)
                y1[b] = make_float2(v.x - w.y, v.y + w.x);
                y2[b] = make_float2(v.x + w.y, v.y - w.x);
            }
            if (c1 >= 0) *reinterpret_cast<float4 *>(ch_iq + (coff + c1) * iq_stride + M) = make_float4(y1[0].x, y1[0].y, y1[1].x, y1[1].y);
            if (c2 >= 0) *reinterpret_cast<float4 *>(ch_iq + (coff + c2) * iq_stride + M) = make_float4(y2[0].x, y2[0].y, y2[1].x, y2[1].y);
        }
    }
}

// ---- host side
struct TableGemm {
    bool enabled = false;
    int D = 0, Lh = 0, N = 0;          // total decimation, prototype length, table period
    int slots = 0, slots_pad = 0, n_pairs = 0;
    int Mpad = 0, n_streams = 1, chans = 0;
    float *d_hc = nullptr;
    float4 *d_tdup = nullptr;          // [N][slots_pad]
    int *d_slot_chan = nullptr;        // [slots_pad][2]
    float2 *d_U = nullptr;             // [n_streams][N][Mpad]
    std::vector<double> h_hc64;        // the composite prototype in double (the tensor-core variant folds with it)
    std::vector<int> h_slot_chan;      // host copy of d_slot_chan (the tensor-core variant lays its rows out from it)
    void destroy() {
        cudaFree(d_hc); cudaFree(d_tdup); cudaFree(d_slot_chan); cudaFree(d_U);
        d_hc = nullptr; d_tdup = nullptr; d_slot_chan = nullptr; d_U = nullptr; enabled = false;
    }
};

// tables: [C][N] float2 of ONE stream (row ignored where has[c] == 0: the table is the constant (1, 0)); partner[] from
// fast_find_partners (bank_fast.cuh) or all -1.
inline int tgemm_init(TableGemm &tg, int C, int N, int n_stages, const int *m, const int *H, const float *taps, const int *tap_off,
                      const float *tables, const uint8_t *has, const std::vector<int> &partner, int max_out, int n_streams) {
    tg.N = N; tg.chans = C; tg.n_streams = n_streams;
    std::vector<double> hc(1, 1.0);
    long up = 1;
    for (int st = 0; st < n_stages; ++st) {                // h_c(z) = prod_s h_s(z^(m_0 .. m_(s-1))), in double
        std::vector<double> nx((hc.size() - 1) + (size_t)(H[st] - 1) * up + 1, 0.0);
        for (size_t i = 0; i < hc.size(); ++i)
            for (int j = 0; j < H[st]; ++j) nx[i + (size_t)j * up] += hc[i] * (double)taps[tap_off[st] + j];
        hc.swap(nx);
        up *= m[st];
    }
    tg.D = (int)up; tg.Lh = (int)hc.size();
    std::vector<float> hcf(hc.begin(), hc.end());
    tg.h_hc64 = hc;
    std::vector<int> slot_chan;
    for (int c = 0; c < C; ++c) {
        if (partner[c] >= 0 && partner[c] < c) continue;   // rides in its partner's slot
        slot_chan.push_back(c);
        slot_chan.push_back(partner[c]);
        if (partner[c] >= 0) ++tg.n_pairs;
    }
    tg.slots = (int)slot_chan.size() / 2;
    tg.slots_pad = (tg.slots + kGemmSlots - 1) / kGemmSlots * kGemmSlots;
    slot_chan.resize((size_t)tg.slots_pad * 2, -1);
    std::vector<float4> tdup((size_t)N * tg.slots_pad, make_float4(0.f, 0.f, 0.f, 0.f));
    for (int s = 0; s < tg.slots; ++s) {
        const int c = slot_chan[2 * s];
        for (int p = 0; p < N; ++p) {
            float tr = 1.0f, ti = 0.0f;                    // no translator: the mix is the identity (msd.hpp:79-101)
            if (tables && has[c]) { tr = tables[((size_t)c * N + p) * 2]; ti = tables[((size_t)c * N + p) * 2 + 1]; }
            tdup[(size_t)p * tg.slots_pad + s] = make_float4(tr, tr, ti, ti);
        }
    }
    tg.h_slot_chan = slot_chan;
    tg.Mpad = (max_out + kGemmM - 1) / kGemmM * kGemmM;
    CU(cudaMalloc((void **)&tg.d_hc, sizeof(float) * hcf.size()));
    CU(cudaMemcpy(tg.d_hc, hcf.data(), sizeof(float) * hcf.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&tg.d_tdup, sizeof(float4) * tdup.size()));
    CU(cudaMemcpy(tg.d_tdup, tdup.data(), sizeof(float4) * tdup.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&tg.d_slot_chan, sizeof(int) * slot_chan.size()));
    CU(cudaMemcpy(tg.d_slot_chan, slot_chan.data(), sizeof(int) * slot_chan.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&tg.d_U, sizeof(float2) * (size_t)n_streams * N * tg.Mpad));
    CU(cudaMemset(tg.d_U, 0, sizeof(float2) * (size_t)n_streams * N * tg.Mpad));
    tg.enabled = true;
    return SDRX_OK;
}

inline bool tgemm_eligible(int N, int n_out) { return N >= kGemmK && N % kGemmK == 0 && n_out > 0 && n_out % 2 == 0; }

template <int FMT>
int tgemm_launch(TableGemm &tg, cudaStream_t st, const void *raw, long stream_bytes, int hist, int valid_lo, int pos0, int n_out,
                 float2 *ch_iq, long iq_stride, uint64_t *launches, cudaEvent_t ev0 = nullptr, cudaEvent_t ev1 = nullptr) {
    k_fold<FMT><<<dim3((n_out + kFoldM - 1) / kFoldM, tg.n_streams), kFoldThreads, 0, st>>>(raw, stream_bytes, hist, valid_lo, pos0, tg.D, tg.N,
                                                                                           tg.d_hc, tg.Lh, n_out, tg.d_U, tg.Mpad);
    if (ev0) CU(cudaEventRecord(ev0, st));
    k_tgemm<<<dim3(tg.slots_pad / kGemmSlots, (n_out + kGemmM - 1) / kGemmM, tg.n_streams), kGemmThreads, kGemmSmem, st>>>(
        tg.d_tdup, tg.slots_pad, tg.d_U, tg.Mpad, tg.N, tg.d_slot_chan, tg.chans, ch_iq, iq_stride, n_out);
    if (ev1) CU(cudaEventRecord(ev1, st));
    *launches += 2;
    CU(cudaGetLastError());
    return SDRX_OK;
}

}  // namespace sdrx

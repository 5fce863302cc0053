// sdrx_b200 -- specialised channel bank for sm_100a: raw->float ingest, NCO mix (msd.hpp:134-135) and
// the first three decimating FIR stages (msd.hpp:251-335) of every channel in ONE kernel, for chains
// whose first three stages all decimate by 5 (the 10 MS/s and 6 MS/s Airspy chains, sdrx.cpp:1646-1668).
//
// Mapping (DESIGN.md "channel bank"):
//   * FP32 CUDA-core bound (19.2 flop per channel-sample against 8 B of *shared* input), so the job is
//     to keep the FMA pipe issuing: accumulations are packed FFMA2 (fma.rn.f32x2 on (re,im) pairs with
//     the tap duplicated), which halves their issue slots and leaves room for the shared-memory loads.
//   * A CTA owns a group of 16 channels.  Their translator tables (the reference's own FP32 tables,
//     full period N, 9.6 KB per channel at N=1200) sit in shared memory for the CTA's whole life; the
//     16 lanes of a half warp are the 16 channels, so one table row is one conflict-free 128 B read.
//   * Time is cut into per-thread segments that are multiples of 125 samples (one stage-3 output).  The
//     two half warps of a warp are two segments; the 16 segments of a CTA stream their input through a
//     double-buffered shared-memory window (global loads for step t+1 are in flight while step t is
//     computed), every sample being fetched from L2 once per 16 channels.
//   * Each thread runs the three stages in transposed form (one accumulator per output in flight), so
//     there are no delay lines: a segment warms up on the 125 samples before it (stages 1-2 recomputed,
//     1.5 % extra work), and stage-3 outputs that straddle a segment boundary are produced as two
//     partial sums (the tail of one segment, the head of the next) that k_fix_tails adds.  At the start
//     of a block the warm-up is 500 samples of the carried raw history and the heads are complete, so
//     the only state carried between blocks is that raw history.
//   * Work is (channel group x time) linearised and cut into one contiguous range per SM, so all 148
//     SMs finish together whatever the channel count.
#pragma once
#include "common.cuh"

namespace sdrx {

constexpr int kFastCh = 16;        // channels per group
constexpr int kFastSegs = 16;      // time segments per CTA
constexpr int kFastThreads = kFastCh * kFastSegs;
constexpr int kUnit = 125;         // raw samples per stage-3 output
constexpr int kBody = 25;          // raw samples per stage-2 output
constexpr int kXStride = 126;      // float2 per staged segment chunk (125 + 1: 1008 B keeps the two
                                   // segments of a warp on different banks)
constexpr int kWrap = 24;          // table rows replicated past N so a 25-sample body never wraps
constexpr int kFastMinSeg = 8;     // units; segments shorter than 4 would overlap their fix-ups
constexpr int kFastWarmStart = 4;  // warm-up units of a segment at the start of the block
constexpr int kFastRawHist = 768;  // raw history samples the engine keeps for this path

typedef unsigned long long u64;

struct FastTaps {                  // kernel parameter: lives in the constant bank
    float2 h1[17];                 // (h,h) pairs for FFMA2
    float2 h2[19];
    float2 h3[24];                 // 21 used
};

struct FastPiece {                 // one (group, unit range) run of one CTA
    int group, u_begin, u_end, nseg, slot_base;
};
struct FastCta { int piece_begin, piece_end; };
struct FastBound { int group, u0, slot; };      // stage-3 outputs u0..u0+3 of `group` need tails[slot]

// the segment split of a piece; shared by the kernel and the host-side schedule builder
__host__ __device__ inline void fast_segment(int u_begin, int u_end, int nseg, int s, int *u0, int *u1) {
    const int len = u_end - u_begin, base = len / nseg, rem = len % nseg;
    *u0 = u_begin + s * base + (s < rem ? s : rem);
    *u1 = *u0 + base + (s < rem ? 1 : 0);
}

__device__ __forceinline__ u64 pack2(float x, float y) { u64 d; asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(x), "f"(y)); return d; }
__device__ __forceinline__ float2 unpack2(u64 v) { float2 d; asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(v)); return d; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ u64 tap2(const float2 &h) { return pack2(h.x, h.y); }

// One 25-sample body of one thread: mixes 25 samples and advances the three stages by 5 / 1 / 0.2 outputs.
//   xp   the thread's 25 staged input samples (shared memory, broadcast within the half warp)
//   tp   table row of the first sample, already offset to the thread's channel (row stride 16 float2)
//   a    stage-1 accumulators: a[0..2] enter in flight, a[5..7] leave in flight (renamed by the caller)
//   b    stage-2 accumulators: b[0..2] enter in flight, b[0] completes here, b[3] starts here
// Stage-1 output 5B+q covers raw 25B + [5q-12, 5q+4]; stage-2 output B+q2 covers stage-1 5B + [5q2-14, 5q2+4].
__device__ __forceinline__ void fast_body(const float2 *__restrict__ xp, const float2 *__restrict__ tp,
                                          const FastTaps &tp_, u64 (&a)[8], u64 (&b)[4]) {
#pragma unroll
    for (int i = 0; i < kBody; ++i) {
        const float2 x = xp[i];
        const float2 w = tp[i * kFastCh];
        const float sr = fmaf(x.x, w.x, -(x.y * w.y));          // msd.hpp:134  in[i] * translator_[pos]
        const float si = fmaf(x.x, w.y, x.y * w.x);
        const u64 s = pack2(sr, si);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int t = i - 5 * q + 12;                         // tap index of this sample in output q
            if (t == 0) a[q] = mul2(s, tap2(tp_.h1[0]));
            else if (t > 0 && t <= 16) a[q] = fma2(s, tap2(tp_.h1[t]), a[q]);
        }
        if (i % 5 == 4) {                                         // stage-1 output q1 = i/5 is complete
            const int q1 = i / 5;
            const u64 y = a[q1];
#pragma unroll
            for (int q2 = 0; q2 < 4; ++q2) {
                const int t2 = q1 - 5 * q2 + 14;
                if (t2 == 0) b[q2] = mul2(y, tap2(tp_.h2[0]));
                else if (t2 > 0 && t2 <= 18) b[q2] = fma2(y, tap2(tp_.h2[t2]), b[q2]);
            }
        }
    }
}

// grid = number of CTAs in the schedule, block = 256, dynamic smem = fast_smem_bytes(N).
template <int FMT>
static __global__ void __launch_bounds__(kFastThreads, 1)
k_bank_fast(const void *__restrict__ raw, int hist, int valid_lo, const float2 *__restrict__ table, int Cpad, int C,
            int N, int pos0, const FastTaps taps, const FastCta *__restrict__ ctas, const FastPiece *__restrict__ pieces,
            int U, float2 *__restrict__ s3, long s3_stride, int s3_off, float2 *__restrict__ tails, int n_slots) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 *tb = reinterpret_cast<float2 *>(smem_raw);                       // [(N + kWrap)][16]
    float2 *xs = tb + (size_t)(N + kWrap) * kFastCh;                         // [2][16][kXStride]
    int *seg_start = reinterpret_cast<int *>(xs + 2 * kFastSegs * kXStride); // [16] first unit staged (may be < 0)
    int *seg_steps = seg_start + kFastSegs;                                  // [16] steps (units) to stage

    const int tid = threadIdx.x;
    const int ch = tid & (kFastCh - 1);
    const int seg = tid >> 4;
    const FastCta me = ctas[blockIdx.x];
    int cur_group = -1;

    for (int pi = me.piece_begin; pi < me.piece_end; ++pi) {
        const FastPiece pc = pieces[pi];
        __syncthreads();                                   // previous piece is done with tb / xs / seg_*
        if (pc.group != cur_group) {
            cur_group = pc.group;
            const float2 *src = table + pc.group * kFastCh;
            for (int idx = tid; idx < (N + kWrap) * kFastCh; idx += kFastThreads) {
                const int row = idx >> 4, c16 = idx & 15;
                tb[idx] = __ldg(src + (long)(row < N ? row : row - N) * Cpad + c16);
            }
        }
        // my segment
        int u0 = 0, u1 = 0, warm = 0;
        const bool have = seg < pc.nseg;
        if (have) {
            fast_segment(pc.u_begin, pc.u_end, pc.nseg, seg, &u0, &u1);
            warm = (u0 == 0) ? kFastWarmStart : 1;
        }
        const int my_steps = have ? warm + (u1 - u0) : 0;
        if (ch == 0) { seg_start[seg] = u0 - warm; seg_steps[seg] = my_steps; }
        int nsteps;
        {
            int a0, a1;
            fast_segment(pc.u_begin, pc.u_end, pc.nseg, 0, &a0, &a1);
            nsteps = ((pc.u_begin == 0) ? kFastWarmStart : 1) + (a1 - a0);   // segment 0 is never the shorter one
        }
        __syncthreads();

        // stage the first chunk
        float2 pre[8];
        auto fetch = [&](int t) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int e = tid + j * kFastThreads;
                const int s = e / kUnit, i = e - s * kUnit;
                pre[j] = make_float2(0.0f, 0.0f);
                if (e < kFastSegs * kUnit && t < seg_steps[s])
                    pre[j] = load_iq_from<FMT>(raw, (long)hist + (long)kUnit * (seg_start[s] + t) + i, valid_lo);
            }
        };
        auto stash = [&](int buf) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int e = tid + j * kFastThreads;
                const int s = e / kUnit, i = e - s * kUnit;
                if (e < kFastSegs * kUnit) xs[(buf * kFastSegs + s) * kXStride + i] = pre[j];
            }
        };
        fetch(0);
        stash(0);
        __syncthreads();

        // per-thread filter state
        u64 a[8], b[4], c3[5];
#pragma unroll
        for (int q = 0; q < 8; ++q) a[q] = 0ull;
#pragma unroll
        for (int q = 0; q < 4; ++q) b[q] = 0ull;
#pragma unroll
        for (int q = 0; q < 5; ++q) c3[q] = 0ull;
        int row = (int)((((long)pos0 + (long)kUnit * (u0 - warm)) % N + N) % N);
        const int feed_from = (u0 == 0) ? 4 : 5 * warm;      // first local body whose stage-2 output feeds stage 3
        const long out_base = (long)(pc.group * kFastCh + ch) * s3_stride + s3_off;
        const bool ch_ok = pc.group * kFastCh + ch < C;

        for (int t = 0; t < nsteps; ++t) {
            if (t + 1 < nsteps) fetch(t + 1);
            if (t < my_steps) {
                const float2 *xseg = xs + ((t & 1) * kFastSegs + seg) * kXStride;
#pragma unroll 1
                for (int p = 0; p < 5; ++p) {
                    fast_body(xseg + p * kBody, tb + row * kFastCh + ch, taps, a, b);
                    row += kBody;
                    if (row >= N) row -= N;
                    const u64 z = b[0];                                  // stage-2 output of this body
                    b[0] = b[1]; b[1] = b[2]; b[2] = b[3];
                    a[0] = a[5]; a[1] = a[6]; a[2] = a[7];
                    // stage 3: output U_+q3 covers stage-2 5U_ + [5q3-16, 5q3+4]
                    if (t * 5 + p >= feed_from) {
                        c3[0] = fma2(z, tap2(taps.h3[p + 16]), c3[0]);
                        c3[1] = fma2(z, tap2(taps.h3[p + 11]), c3[1]);
                        c3[2] = fma2(z, tap2(taps.h3[p + 6]), c3[2]);
                        c3[3] = fma2(z, tap2(taps.h3[p + 1]), c3[3]);
                        if (p == 4) c3[4] = mul2(z, tap2(taps.h3[0]));
                    }
                }
                if (t >= warm && ch_ok) s3[out_base + (u0 + t - warm)] = unpack2(c3[0]);
                c3[0] = c3[1]; c3[1] = c3[2]; c3[2] = c3[3]; c3[3] = c3[4]; c3[4] = 0ull;
            }
            if (t + 1 < nsteps) stash((t + 1) & 1);
            __syncthreads();
        }
        // what this segment contributes to the first four outputs of the next one
        if (have && u1 < U && ch_ok) {
            float2 *tl = tails + ((long)(pc.group * kFastCh + ch) * n_slots + pc.slot_base + seg) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) tl[q] = unpack2(c3[q]);
        }
    }
}

// s3[c][u0 + i] += tails[c][slot][i]: completes the stage-3 outputs that straddle a segment boundary.
static __global__ void k_fix_tails(float2 *__restrict__ s3, long s3_stride, int s3_off, const float2 *__restrict__ tails,
                                   int n_slots, const FastBound *__restrict__ bounds, int n_bounds, int C) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_bounds * kFastCh * 4) return;
    const int i = idx & 3, c16 = (idx >> 2) & 15, bi = idx >> 6;
    const FastBound bd = bounds[bi];
    const int c = bd.group * kFastCh + c16;
    if (c >= C) return;
    const float2 t = tails[((long)c * n_slots + bd.slot) * 4 + i];
    float2 *p = s3 + (long)c * s3_stride + s3_off + bd.u0 + i;
    float2 v = *p;
    v.x += t.x; v.y += t.y;
    *p = v;
}

inline size_t fast_smem_bytes(int N) {
    return sizeof(float2) * ((size_t)(N + kWrap) * kFastCh + 2 * kFastSegs * kXStride) + sizeof(int) * 2 * kFastSegs;
}

// ------------------------------------------------------------------------------------------------
// Host side: the static schedule (depends only on C, the block length and the SM count) and launch.
struct FastSchedule {
    int U = 0;                         // units (stage-3 outputs per channel) of the block length it was built for
    int n_ctas = 0, n_slots = 0, n_bounds = 0;
    FastCta *d_ctas = nullptr;
    FastPiece *d_pieces = nullptr;
    FastBound *d_bounds = nullptr;
};

struct FastBank {
    bool enabled = false;
    int stages_fused = 0;              // leading MSD stages the kernel covers (3)
    int n_sms = 148;
    int groups = 0;
    FastTaps taps{};
    FastSchedule sched[2];             // small cache keyed by U (block length / 125)
    float2 *d_tails = nullptr;
    int tails_slots = 0;               // capacity of d_tails in slots per channel
    int C = 0;

    void free_schedule(FastSchedule &s) {
        cudaFree(s.d_ctas); cudaFree(s.d_pieces); cudaFree(s.d_bounds);
        s = FastSchedule{};
    }
    void destroy() {
        for (auto &s : sched) free_schedule(s);
        cudaFree(d_tails);
        d_tails = nullptr;
        enabled = false;
    }
};

// Which chains the kernel covers: >= 3 stages, the first three decimating by 5 with 17/19/21 taps.
inline bool fast_eligible(int S, const int *m, const int *H, int N) {
    if (S < 3) return false;
    if (m[0] != 5 || m[1] != 5 || m[2] != 5) return false;
    if (H[0] != 17 || H[1] != 19 || H[2] != 21) return false;
    if (N < 1 || fast_smem_bytes(N) > 227 * 1024) return false;
    return true;
}

// Linearise (group, unit) work, cut it into one contiguous range per CTA, snap cuts that fall close to
// a group boundary onto it, split ranges at group boundaries into pieces, and number the segment-end
// boundaries whose stage-3 partial sums k_fix_tails has to add.
inline int fast_build_schedule(FastBank &fb, FastSchedule &out, int U) {
    std::vector<FastCta> ctas;
    std::vector<FastPiece> pieces;
    std::vector<FastBound> bounds;
    const long total = (long)fb.groups * U;
    const long min_range = (long)kFastSegs * kFastMinSeg;                // 128 units: every segment >= 8 units
    long n_ctas = std::min<long>(fb.n_sms, std::max<long>(1, total / min_range));
    const int snap = std::min(4 * kFastMinSeg, std::max(1, U / 2));
    std::vector<long> cuts;
    cuts.push_back(0);
    for (long i = 1; i < n_ctas; ++i) {
        long c = (total * i) / n_ctas;
        const long g = c / U, u = c % U;
        if (u < snap) c = g * U;
        else if (U - u < snap) c = (g + 1) * U;
        if (c > cuts.back() && c < total) cuts.push_back(c);
    }
    cuts.push_back(total);
    int slot = 0, slot_group = -1, max_slots = 0;     // slots are numbered per group (pieces of a group are consecutive)
    for (size_t i = 0; i + 1 < cuts.size(); ++i) {
        FastCta cta{(int)pieces.size(), 0};
        long w = cuts[i];
        while (w < cuts[i + 1]) {
            const int g = (int)(w / U), ub = (int)(w % U);
            const int ue = (int)std::min<long>(U, ub + (cuts[i + 1] - w));
            if (g != slot_group) { slot_group = g; slot = 0; }
            FastPiece p{g, ub, ue, 0, slot};
            p.nseg = std::max(1, std::min(kFastSegs, (ue - ub) / kFastMinSeg));
            for (int s = 0; s < p.nseg; ++s) {
                int u0, u1;
                fast_segment(ub, ue, p.nseg, s, &u0, &u1);
                if (u1 < U) bounds.push_back(FastBound{g, u1, slot + s});
            }
            slot += p.nseg;
            max_slots = std::max(max_slots, slot);
            pieces.push_back(p);
            w += ue - ub;
        }
        cta.piece_end = (int)pieces.size();
        ctas.push_back(cta);
    }
    out.U = U;
    out.n_ctas = (int)ctas.size();
    out.n_slots = max_slots;           // tails are indexed [channel][slot][4]
    out.n_bounds = (int)bounds.size();
    CU(cudaMalloc((void **)&out.d_ctas, sizeof(FastCta) * ctas.size()));
    CU(cudaMalloc((void **)&out.d_pieces, sizeof(FastPiece) * pieces.size()));
    CU(cudaMalloc((void **)&out.d_bounds, sizeof(FastBound) * std::max<size_t>(1, bounds.size())));
    CU(cudaMemcpy(out.d_ctas, ctas.data(), sizeof(FastCta) * ctas.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(out.d_pieces, pieces.data(), sizeof(FastPiece) * pieces.size(), cudaMemcpyHostToDevice));
    if (!bounds.empty())
        CU(cudaMemcpy(out.d_bounds, bounds.data(), sizeof(FastBound) * bounds.size(), cudaMemcpyHostToDevice));
    return SDRX_OK;
}

// Launch the bank for one block of n = 125*U samples (all stage phases zero), then the tail fix-up.
template <int FMT>
int fast_launch(FastBank &fb, cudaStream_t st, const void *raw, int hist, int valid_lo, const float2 *table, int Cpad,
                int N, int pos0, unsigned n, float2 *s3, long s3_stride, int s3_off, uint64_t *launches) {
    const int U = (int)(n / kUnit);
    FastSchedule *sc = nullptr;
    for (auto &s : fb.sched) if (s.U == U) sc = &s;
    if (!sc) {
        FastSchedule &victim = fb.sched[0].U == 0 ? fb.sched[0] : fb.sched[1];
        CU(cudaStreamSynchronize(st));          // a launch may still be reading the evicted schedule
        fb.free_schedule(victim);
        int rc = fast_build_schedule(fb, victim, U);
        if (rc) return rc;
        sc = &victim;
    }
    if (sc->n_slots > fb.tails_slots) {
        CU(cudaStreamSynchronize(st));
        cudaFree(fb.d_tails);
        fb.d_tails = nullptr;
        CU(cudaMalloc((void **)&fb.d_tails, sizeof(float2) * 4 * (size_t)sc->n_slots * fb.C));
        fb.tails_slots = sc->n_slots;
    }
    k_bank_fast<FMT><<<sc->n_ctas, kFastThreads, fast_smem_bytes(N), st>>>(
        raw, hist, valid_lo, table, Cpad, fb.C, N, pos0, fb.taps, sc->d_ctas, sc->d_pieces, U, s3, s3_stride, s3_off,
        fb.d_tails, fb.tails_slots);
    ++*launches;
    if (sc->n_bounds > 0) {
        const int threads = sc->n_bounds * kFastCh * 4;
        k_fix_tails<<<div_up(threads, 256), 256, 0, st>>>(s3, s3_stride, s3_off, fb.d_tails, fb.tails_slots, sc->d_bounds,
                                                          sc->n_bounds, fb.C);
        ++*launches;
    }
    CU(cudaGetLastError());
    return SDRX_OK;
}

inline int fast_init(FastBank &fb, int C, int N, const float *h1, const float *h2, const float *h3) {
    fb.C = C;
    fb.groups = (C + kFastCh - 1) / kFastCh;
    fb.stages_fused = 3;
    for (int i = 0; i < 17; ++i) fb.taps.h1[i] = make_float2(h1[i], h1[i]);
    for (int i = 0; i < 19; ++i) fb.taps.h2[i] = make_float2(h2[i], h2[i]);
    for (int i = 0; i < 24; ++i) fb.taps.h3[i] = i < 21 ? make_float2(h3[i], h3[i]) : make_float2(0.0f, 0.0f);
    int dev = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&fb.n_sms, cudaDevAttrMultiProcessorCount, dev));
    const int smem = (int)fast_smem_bytes(N);
    CU(cudaFuncSetAttribute(k_bank_fast<SDRX_FMT_U8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CU(cudaFuncSetAttribute(k_bank_fast<SDRX_FMT_S16>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    CU(cudaFuncSetAttribute(k_bank_fast<SDRX_FMT_F32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    fb.enabled = true;
    return SDRX_OK;
}

}  // namespace sdrx

// sdrx_b200 -- specialised channel bank for sm_100a: raw->float ingest, NCO mix (msd.hpp:134-135) and
// the first three decimating FIR stages (msd.hpp:251-335) of every channel in ONE kernel, for chains
// whose first three stages all decimate by 5 (the 10 MS/s Airspy chain, sdrx.cpp:1660-1668).
//
// Mapping (DESIGN.md "channel bank"):
//   * FP32 CUDA-core bound (19.2 flop per channel-sample against 8 B of *shared* input), so the job is
//     to keep the FMA pipe issuing: accumulations are packed FFMA2 (fma.rn.f32x2 on (re,im) pairs with
//     the tap duplicated), which halves their issue slots and leaves room for the shared-memory loads.
//   * A CTA owns a group of 16 lane slots.  Their translator tables (the reference's own FP32 tables,
//     full period N, 9.6 KB per slot at N=1200) sit in shared memory for the CTA's whole life; the
//     16 lanes of a half warp are the 16 slots, so one table row is one conflict-free 128 B read.
//   * NCH = 2 ("conjugate pairs"): the table of channel offset -k is, bit for bit, the complex conjugate of
//     the table of +k (the float-rounded angle is sign symmetric, cosf is even, sinf odd), so one thread
//     carries BOTH channels of a +-k pair off one table word and one input word: the shared-memory traffic per
//     channel-sample halves and the two mixes share their four products (6 instead of 8 FMA-pipe slots).
//     The host pairs channels by comparing the tables it was given (never by assuming offsets); channels
//     without a partner ride alone in a slot.  NCH = 1 is the same kernel with one channel per slot and is
//     used when too few channels pair up.
//   * Time is cut into per-thread segments that are multiples of 125 samples (one stage-3 output).  The
//     two half warps of a warp are two segments; each segment streams its input through its own
//     double-buffered shared-memory window (global loads for step t+1 are in flight while step t is
//     computed; the window is private to the half warp, so only warp-level syncs are needed), every
//     sample being fetched from L2 once per slot group.
//   * Each thread runs the three stages in transposed form (one accumulator per output in flight), so
//     there are no delay lines: a segment warms up on the 125 samples before it (stages 1-2 recomputed),
//     and stage-3 outputs that straddle a segment boundary are produced as two partial sums (the tail of
//     one segment, the head of the next) that the next kernel adds in a fixed order.  At the start
//     of a block the warm-up is 500 samples of the carried raw history and the heads are complete, so
//     the only state carried between blocks is that raw history.
//   * Work is (slot group x time) linearised and cut into one contiguous range per SM, so all 148
//     SMs finish together whatever the channel count.
#pragma once
#include <algorithm>
#include <cstring>
#include <map>
#include <vector>

#include "common.cuh"
#include "host_common.h"

namespace sdrx {

constexpr int kFastCh = 16;        // lane slots per group
constexpr int kFastSegs = 32;      // time segments per CTA (512 threads: 4 warps per scheduler)
constexpr int kFastThreads = kFastCh * kFastSegs;
constexpr int kUnit = 125;         // raw samples per stage-3 output
constexpr int kBody = 25;          // raw samples per stage-2 output
constexpr int kXBody = 26;         // float2 per staged 25-sample body (+1: every body starts 16 B aligned)
constexpr int kXRing = 3;          // staged bodies per segment: the one being computed and two in flight (a body takes a
                                   // 16-warp SM ~1.5 us, an L2 round trip ~1 us)
constexpr int kXSeg = kXRing * kXBody;   // float2 per segment's ring; 624 B puts the two segments of a warp 28 banks apart,
                                   // so their 16 B broadcast reads never collide
constexpr int kWrap = 24;          // table rows replicated past N so a 25-sample body never wraps
constexpr int kFastMinSeg = 8;     // units per segment when sizing the grid: shorter ones mostly warm up
constexpr int kFastMinSegPiece = 2;// ... but a small piece (a CTA range that straddles a group) still spreads over all warps
constexpr int kFastWarmStart = 4;  // warm-up units of a segment at the start of the block
constexpr int kFastPiecePenalty = 1;// steps charged to a CTA for every piece after its first (table reload + refill)
constexpr int kFastSideSms = 3;    // SMs the schedule leaves without a bank CTA (== kSideSms, demod.cuh)
constexpr int kFastRawHist = 768;  // raw history samples the engine keeps for this path

struct FastTaps {                  // kernel parameter: lives in the constant bank
    float2 h1[17];                 // (h,h) pairs for FFMA2
    float2 h2[19];
    float2 h3[24];                 // 21 used
};

// One (group, unit range) run of one CTA.  The range is cut into two halves A = [u_begin, u_begin+H) and
// B = [u_begin+H, u_end); warp w takes the w-th segment (seg_len units) of A on lanes 0-15 and the w-th
// segment of B on lanes 16-31.
struct FastPiece {
    int group, u_begin, u_end, H, seg_len, slot_base;
};
struct FastCta { int piece_begin, piece_end; };
typedef TailFix FastFix;                        // stage-3 output u of `group` += tails[src[0..n)]
constexpr int kFastSlotsPerPiece = 2 * (kFastSegs / 2);

// The segment that starts a block warms up on kFastWarmStart units of raw history instead of one; it is made
// that much shorter so that every segment of the piece runs the same number of steps.
__host__ __device__ inline int fast_cut(const FastPiece &p) {
    return (p.u_begin == 0 && p.seg_len > 2 * (kFastWarmStart - 1)) ? kFastWarmStart - 1 : 0;
}
// the segment of (warp w, half) in a piece; shared by the kernel and the host-side schedule builder
__host__ __device__ inline bool fast_segment(const FastPiece &p, int w, int half, int *u0, int *u1) {
    const int lo = half ? p.u_begin + p.H : p.u_begin;
    const int hi = half ? p.u_end : p.u_begin + p.H;
    const int cut = half ? 0 : fast_cut(p);
    *u0 = lo + w * p.seg_len - (w > 0 ? cut : 0);
    const int e = *u0 + p.seg_len - (w == 0 ? cut : 0);
    *u1 = e < hi ? e : hi;
    return *u0 < hi;
}

__device__ __forceinline__ u64 tap2(const float2 &h) { return pack2(h.x, h.y); }

// One 25-sample body of one thread: mixes 25 samples and advances the three stages by 5 / 1 / 0.2 outputs, for
// the NCH channels of its slot (channel 1 mixes with the conjugate of the table word).
//   xp   the thread's 25 staged input samples (shared memory, broadcast within the half warp)
//   tp   table row of the first sample, already offset to the thread's slot (row stride 16 float2)
//   a    stage-1 accumulators: a[0..2] enter in flight, a[5..7] leave in flight (renamed by the caller)
//   b    stage-2 accumulators: b[0..2] enter in flight, b[0] completes here, b[3] starts here
// Stage-1 output 5B+q covers raw 25B + [5q-12, 5q+4]; stage-2 output B+q2 covers stage-1 5B + [5q2-14, 5q2+4].
template <int NCH>
__device__ __forceinline__ void fast_body(const float2 *__restrict__ xp, const float2 *__restrict__ tp,
                                          const FastTaps &tp_, u64 (&a)[NCH][8], u64 (&b)[NCH][4]) {
    float2 xv[kBody + 1];
#pragma unroll
    for (int i = 0; i < kBody + 1; i += 2) {                   // 13 x LDS.128 (the 26th slot is padding)
        const float4 v = *reinterpret_cast<const float4 *>(xp + i);
        xv[i] = make_float2(v.x, v.y);
        xv[i + 1] = make_float2(v.z, v.w);
    }
#pragma unroll
    for (int i = 0; i < kBody; ++i) {
        const float2 x = xv[i];
        const float2 w = tp[i * kFastCh];
        // msd.hpp:134  in[i] * translator_[pos].  A conjugate pair does not form the two mixed samples at all: the
        // stages are linear, so the four real product streams P = xr wr, S = xi wr, R = xr wi, Q = xi wi are
        // filtered as the packed pairs (P,S) and (R,Q) -- 4 multiplies per sample for both channels -- and
        // x w = (P - Q, R + S), x conj(w) = (P + Q, S - R) are taken on the stage-3 results (fast_unmix).
        u64 s[NCH];
        if (NCH == 2) {
            s[0] = pack2(x.x * w.x, x.y * w.x);
            s[1] = pack2(x.x * w.y, x.y * w.y);
        } else {
            const float m1 = x.y * w.y, m2 = x.y * w.x;
            s[0] = pack2(fmaf(x.x, w.x, -m1), fmaf(x.x, w.y, m2));
        }
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
                const int t = i - 5 * q + 12;                     // tap index of this sample in output q
                if (t == 0) a[c][q] = mul2(s[c], tap2(tp_.h1[0]));
                else if (t > 0 && t <= 16) a[c][q] = fma2(s[c], tap2(tp_.h1[t]), a[c][q]);
            }
            if (i % 5 == 4) {                                     // stage-1 output q1 = i/5 is complete
                const int q1 = i / 5;
                const u64 y = a[c][q1];
#pragma unroll
                for (int q2 = 0; q2 < 4; ++q2) {
                    const int t2 = q1 - 5 * q2 + 14;
                    if (t2 == 0) b[c][q2] = mul2(y, tap2(tp_.h2[0]));
                    else if (t2 > 0 && t2 <= 18) b[c][q2] = fma2(y, tap2(tp_.h2[t2]), b[c][q2]);
                }
            }
        }
    }
}

// Stage-3 sums of the product streams (P,S) and (R,Q) of a conjugate pair -> the pair's two channel samples.
__device__ __forceinline__ void fast_unmix(u64 ps, u64 rq, float2 &plus, float2 &minus) {
    const float2 a = unpack2(ps), b = unpack2(rq);
    plus = make_float2(a.x - b.y, b.x + a.y);
    minus = make_float2(a.x + b.y, a.y - b.x);
}

// Debug trace (tools/diag_bank_trace.py): when set, every warp records the globaltimer at its start and end.
__device__ unsigned long long *g_bank_trace = nullptr;
__device__ __forceinline__ unsigned long long bank_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// grid = number of CTAs in the schedule, block = kFastThreads, dynamic smem = fast_smem_bytes(N).
//   table      [N][tstride] float2, one column per lane slot (slot = group * 16 + lane)
//   slot_chan  [slots][NCH] channel index carried by the slot (-1: none); channel 1 of a slot is the conjugate partner
template <int FMT, int NCH>
static __global__ void __launch_bounds__(kFastThreads, 1)
k_bank_fast(const void *__restrict__ raw_all, int hist, int valid_lo, const float2 *__restrict__ table, int tstride,
            const int *__restrict__ slot_chan, int N, int pos0, const FastTaps taps, const FastCta *__restrict__ ctas,
            const FastPiece *__restrict__ pieces, int U, float2 *__restrict__ s3, long s3_stride, int s3_off,
            float2 *__restrict__ tails, int n_slots, int gps, long stream_bytes) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 *tb = reinterpret_cast<float2 *>(smem_raw);                       // [(N + kWrap)][16]
    float2 *xs = tb + (size_t)(N + kWrap) * kFastCh;                         // [kFastSegs][kXRing][kXBody]

    const int tid = threadIdx.x;
    const int ch = tid & (kFastCh - 1);
    const int half = (tid >> 4) & 1;
    const int w = tid >> 5;
    const int seg = tid >> 4;                               // == 2*w + half: slot of the staged chunk
    const FastCta me = ctas[blockIdx.x];
    int cur_group = -1;
    unsigned long long *const trace = g_bank_trace;
    if (trace && (tid & 31) == 0) trace[(blockIdx.x * (kFastThreads / 32) + w) * 2] = bank_now();

    for (int pi = me.piece_begin; pi < me.piece_end; ++pi) {
        const FastPiece pc = pieces[pi];
        // multi-stream engine: slot group g belongs to stream g / gps and uses the translator tables of group g % gps (every
        // stream listens to the same channel plan); the stream's raw block lies stream_bytes further
        const int tgroup = pc.group % gps;
        const void *raw = reinterpret_cast<const unsigned char *>(raw_all) + (long)(pc.group / gps) * stream_bytes;
        if (tgroup != cur_group) {
            __syncthreads();                               // every warp is done with the previous group's table
            cur_group = tgroup;
            const float2 *src = table + tgroup * kFastCh;             // 16 B aligned: tstride and the group base are multiples of 16
            for (int idx = tid; idx < (N + kWrap) * (kFastCh / 2); idx += kFastThreads) {   // all copies in flight at once
                const int row = idx >> 3, c8 = idx & 7;
                const unsigned dst = (unsigned)__cvta_generic_to_shared(tb + row * kFastCh + 2 * c8);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src + (long)(row < N ? row : row - N) * tstride + 2 * c8) : "memory");
            }
            asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
            __syncthreads();
        }
        // my segment
        int u0 = 0, u1 = 0;
        const bool have = fast_segment(pc, w, half, &u0, &u1);
        const int warm = (u0 == 0) ? kFastWarmStart : 1;
        const int my_steps = have ? warm + (u1 - u0) : 0;
        const int nsteps = ((pc.u_begin == 0 && !fast_cut(pc)) ? kFastWarmStart : 1) + pc.seg_len;   // the longest segment's

        // Staging: the 16 threads of a segment fetch that segment's samples one 25-sample body at a time into a ring of
        // kXRing bodies, two bodies ahead of the one being computed.  F32 goes global->shared with cp.async; u8/s16 are
        // loaded early into registers, converted and stored after the compute.  (A whole double-buffered unit per segment
        // used to sit here: 66 KB of the SM; the ring is 20 KB, and what it frees lets the last-stage kernel of the
        // previous block run NEXT TO the bank instead of between two banks.)
        // The hot loop is issue-bound, so the per-body bookkeeping is kept to running offsets: fo / co = byte offset of the
        // ring slot the next fetch fills / the next body reads (0, 208, 416), gidx = raw index of this lane's first sample of
        // the next body to fetch.
        const long my_raw0 = (long)hist + (long)kUnit * (u0 - warm) + ch;
        float2 *my_xs = xs + seg * kXSeg;
        const int my_bodies = 5 * my_steps;
        constexpr unsigned kSlotBytes = kXBody * sizeof(float2);
        unsigned ring_s = (unsigned)__cvta_generic_to_shared(my_xs) + (unsigned)ch * (unsigned)sizeof(float2);
        const bool second = ch < kBody - 16;               // lanes 0-8 also carry samples 16-24 of a body
        unsigned fo = 0, co = 0;
        long gidx = my_raw0;
        int to_fetch = my_bodies;                          // bodies not yet fetched
        // The F32 path keeps its two running addresses opaque to the compiler (it would otherwise rebuild them from the
        // thread index on every fetch to save two registers: ~20 instructions per body in a loop that is issue-bound).
        const unsigned char *gp = reinterpret_cast<const unsigned char *>(reinterpret_cast<const float2 *>(raw) + my_raw0);
        unsigned fs = ring_s;                              // shared-memory address the next fetch writes
        const unsigned ring_end = ring_s + kXRing * kSlotBytes;
        asm volatile("" : "+l"(gp), "+r"(fs), "+r"(ring_s));
        typename RawTraits<FMT>::vec pre[2];
        long pre_idx = 0;
        unsigned pre_off = 0;
        bool pre_valid = false;
        auto fetch = [&]() {                               // the next body of this segment -> its ring slot
            if constexpr (FMT == SDRX_FMT_F32) {
                if (to_fetch > 0) {
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(fs), "l"(gp) : "memory");
                    if (second) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(fs + 128u), "l"(gp + 128) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                --to_fetch;
                gp += kBody * sizeof(float2);
                fs += kSlotBytes;
                if (fs == ring_end) fs = ring_s;
                asm volatile("" : "+l"(gp), "+r"(fs));
                return;
            }
            if (to_fetch > 0) {
                if constexpr (FMT == SDRX_FMT_S16) {
                    unsigned v0, v1 = 0;
                    asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v0) : "l"(reinterpret_cast<const short2 *>(raw) + gidx));
                    if (second) asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v1) : "l"(reinterpret_cast<const short2 *>(raw) + gidx + 16));
                    pre[0] = make_short2((short)(v0 & 0xffff), (short)(v0 >> 16));
                    pre[1] = make_short2((short)(v1 & 0xffff), (short)(v1 >> 16));
                } else if constexpr (FMT == SDRX_FMT_U8) {
                    unsigned short v0, v1 = 0;
                    asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(v0) : "l"(reinterpret_cast<const uchar2 *>(raw) + gidx));
                    if (second) asm volatile("ld.global.nc.u16 %0, [%1];" : "=h"(v1) : "l"(reinterpret_cast<const uchar2 *>(raw) + gidx + 16));
                    pre[0] = make_uchar2((unsigned char)(v0 & 0xff), (unsigned char)(v0 >> 8));
                    pre[1] = make_uchar2((unsigned char)(v1 & 0xff), (unsigned char)(v1 >> 8));
                }
            }
            pre_valid = to_fetch > 0; pre_idx = gidx; pre_off = fo;
            --to_fetch;
            gidx += kBody;
            fo = fo == 2 * kSlotBytes ? 0u : fo + kSlotBytes;
        };
        // make the body fetched last-but-one (F32: all but the youngest cp.async group; u8/s16: what the last fetch left in
        // registers) visible in its slot, before the warp sync
        auto land = [&]() {
            if constexpr (FMT == SDRX_FMT_F32) {
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                if (!pre_valid) return;
                float2 *slot = reinterpret_cast<float2 *>(reinterpret_cast<unsigned char *>(my_xs) + pre_off) + ch;
                float2 v = convert_iq<FMT>(pre[0]);
                if (pre_idx < valid_lo) v = make_float2(0.0f, 0.0f);                // before the stream
                slot[0] = v;
                if (second) {
                    float2 w2 = convert_iq<FMT>(pre[1]);
                    if (pre_idx + 16 < valid_lo) w2 = make_float2(0.0f, 0.0f);
                    slot[16] = w2;
                }
            }
        };
        __syncwarp();                                      // the warp is done with its previous piece's ring
        if constexpr (FMT == SDRX_FMT_F32) {
            fetch();                                       // body 0
            fetch();                                       // body 1
            land();                                        // body 0 is in; body 1 may still be in flight
        } else {
            fetch();
            land();                                        // body 0 is in
            fetch();                                       // body 1 stays in registers until the end of body 0
        }
        __syncwarp();                                      // a segment's ring is private to its half warp: no CTA barrier

        // per-thread filter state
        u64 a[NCH][8], b[NCH][4], c3[NCH][5];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
#pragma unroll
            for (int q = 0; q < 8; ++q) a[c][q] = 0ull;
#pragma unroll
            for (int q = 0; q < 4; ++q) b[c][q] = 0ull;
#pragma unroll
            for (int q = 0; q < 5; ++q) c3[c][q] = 0ull;
        }
        int row = (int)((((long)pos0 + (long)kUnit * (u0 - warm)) % N + N) % N);
        const int feed_from = (u0 == 0) ? 4 : 5 * warm;      // first local body whose stage-2 output feeds stage 3
        int chan[NCH];
        long out_base[NCH];
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            chan[c] = slot_chan[(pc.group * kFastCh + ch) * NCH + c];
            out_base[c] = (long)chan[c] * s3_stride + s3_off;
        }

        for (int t = 0; t < nsteps; ++t) {
#pragma unroll 1
            for (int p = 0; p < 5; ++p) {
                // F32: body jb + 2 goes into the slot body jb - 1 left (the warp sync at the end of jb - 1 has passed).
                // u8/s16: body jb + 1 is in registers (fetched during jb - 1) and is stored at the end of jb.
                if constexpr (FMT == SDRX_FMT_F32) fetch();
                if (t < my_steps) {
                    fast_body<NCH>(reinterpret_cast<const float2 *>(reinterpret_cast<const unsigned char *>(my_xs) + co),
                                   tb + row * kFastCh + ch, taps, a, b);
                    co = co == 2 * kSlotBytes ? 0u : co + kSlotBytes;
                    row += kBody;
                    if (row >= N) row -= N;
#pragma unroll
                    for (int c = 0; c < NCH; ++c) {
                        const u64 z = b[c][0];                               // stage-2 output of this body
                        b[c][0] = b[c][1]; b[c][1] = b[c][2]; b[c][2] = b[c][3];
                        a[c][0] = a[c][5]; a[c][1] = a[c][6]; a[c][2] = a[c][7];
                        // stage 3: output U_+q3 covers stage-2 5U_ + [5q3-16, 5q3+4]
                        if (t * 5 + p >= feed_from) {
                            c3[c][0] = fma2(z, tap2(taps.h3[p + 16]), c3[c][0]);
                            c3[c][1] = fma2(z, tap2(taps.h3[p + 11]), c3[c][1]);
                            c3[c][2] = fma2(z, tap2(taps.h3[p + 6]), c3[c][2]);
                            c3[c][3] = fma2(z, tap2(taps.h3[p + 1]), c3[c][3]);
                            if (p == 4) c3[c][4] = mul2(z, tap2(taps.h3[0]));
                        }
                    }
                }
                land();                                    // F32: at most the youngest group (body jb + 2) still in flight
                if constexpr (FMT != SDRX_FMT_F32) fetch();
                __syncwarp();
            }
            if (t < my_steps) {
                if (t >= warm) {
                    float2 y[2];
                    if (NCH == 2) fast_unmix(c3[0][0], c3[NCH - 1][0], y[0], y[1]);
                    else y[0] = unpack2(c3[0][0]);
#pragma unroll
                    for (int c = 0; c < NCH; ++c)
                        if (chan[c] >= 0) s3[out_base[c] + (u0 + t - warm)] = y[c];
                }
#pragma unroll
                for (int c = 0; c < NCH; ++c) {
                    c3[c][0] = c3[c][1]; c3[c][1] = c3[c][2]; c3[c][2] = c3[c][3]; c3[c][3] = c3[c][4]; c3[c][4] = 0ull;
                }
            }
        }
        if constexpr (FMT == SDRX_FMT_F32) asm volatile("cp.async.wait_group 0;" ::: "memory");
        // what this segment contributes to the (up to four) outputs that follow it
        if (have && u1 < U) {
            float2 y[4][2];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (NCH == 2) fast_unmix(c3[0][q], c3[NCH - 1][q], y[q][0], y[q][1]);
                else y[q][0] = unpack2(c3[0][q]);
            }
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                if (chan[c] < 0) continue;
                float2 *tl = tails + ((long)chan[c] * n_slots + pc.slot_base + half * (kFastSegs / 2) + w) * 4;
#pragma unroll
                for (int q = 0; q < 4; ++q) tl[q] = y[q][c];
            }
        }
    }
    if (trace && (tid & 31) == 0) trace[(blockIdx.x * (kFastThreads / 32) + w) * 2 + 1] = bank_now();
}

// Completes the stage-3 outputs that straddle segment boundaries (only when the next kernel does not do it itself):
// s3[c][u] += sum of the tails that reach u, added in a fixed order by one thread per (output, channel of the group).
static __global__ void k_fix_tails(float2 *__restrict__ s3, long s3_stride, int s3_off, const float2 *__restrict__ tails,
                                   int n_slots, const FastFix *__restrict__ fixes, int n_fixes,
                                   const int *__restrict__ slot_chan, int per_group /* 16 * NCH */) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_fixes * per_group) return;
    const int j = idx % per_group, fi = idx / per_group;
    const FastFix fx = fixes[fi];
    const int c = slot_chan[fx.group * per_group + j];
    if (c < 0) return;
    float2 *p = s3 + (long)c * s3_stride + s3_off + fx.u;
    float2 v = *p;
    for (int k = 0; k < fx.n; ++k) {
        const float2 t = tails[(long)c * n_slots * 4 + fx.src[k]];
        v.x += t.x; v.y += t.y;
    }
    *p = v;
}

inline size_t fast_smem_bytes(int N) {
    return sizeof(float2) * ((size_t)(N + kWrap) * kFastCh + kFastSegs * kXSeg);
}

// ------------------------------------------------------------------------------------------------
// Host side: slot assignment (conjugate pairing), the static schedule (depends only on the slot count,
// the block length and the SM count) and launch.
struct FastSchedule {
    int U = 0;                         // units (stage-3 outputs per channel) of the block length it was built for
    int n_ctas = 0, n_slots = 0, n_fixes = 0;
    FastCta *d_ctas = nullptr;
    FastPiece *d_pieces = nullptr;
    FastFix *d_fixes = nullptr;
    int2 *d_fix_range = nullptr;       // [groups] -> [begin, end) into d_fixes (the list is sorted by group)
};

struct FastBank {
    bool enabled = false;
    int stages_fused = 0;              // leading MSD stages the kernel covers (3)
    int n_sms = 148;
    int nch = 1;                       // channels per lane slot: 2 when conjugate pairing pays
    int groups = 0;                    // slot groups of 16 over ALL streams: n_streams * gps
    int gps = 0;                       // groups per stream (ceil(slots of one stream / 16))
    int n_streams = 1;
    int n_pairs = 0;
    int period_units = 1;              // units after which the translator phase repeats: N / gcd(N, 125)
    FastTaps taps{};
    FastSchedule sched[2];             // small cache keyed by U (block length / 125)
    float2 *d_tails = nullptr;         // two halves (ping-pong) of [C][tails_slots][4] partial stage-3 sums across segment
                                       // boundaries: the last-stage kernel of block b reads one under the bank of block
                                       // b+1, which writes the other
    int tails_slots = 0;               // capacity of one half in slots per channel
    float2 *tails_half(int h) const { return d_tails + (size_t)h * 4 * (size_t)tails_slots * C; }
    int C = 0;
    float2 *d_table_slot = nullptr;    // [N][groups*16] one column per slot (own copy only when paired)
    int tstride = 0;
    int *d_slot_chan = nullptr;        // [groups*16][nch]
    int *d_chan_group = nullptr;       // [C] group whose fix list covers the channel

    void free_schedule(FastSchedule &s) {
        cudaFree(s.d_ctas); cudaFree(s.d_pieces); cudaFree(s.d_fixes); cudaFree(s.d_fix_range);
        s = FastSchedule{};
    }
    void destroy() {
        for (auto &s : sched) free_schedule(s);
        cudaFree(d_tails); cudaFree(d_table_slot); cudaFree(d_slot_chan); cudaFree(d_chan_group);
        d_tails = nullptr; d_table_slot = nullptr; d_slot_chan = nullptr; d_chan_group = nullptr;
        enabled = false;
    }
};

// Which chains the kernel covers: >= 3 stages, the first three decimating by 5 with 17/19/21 taps.
inline bool fast_eligible(int S, const int *m, const int *H, int N) {
    if (S < 3) return false;
    if (m[0] != 5 || m[1] != 5 || m[2] != 5) return false;
    if (H[0] != 17 || H[1] != 19 || H[2] != 21) return false;
    // the table is staged with kWrap rows replicated past N and the row index wraps by one subtraction per 25-sample
    // body: both need N > kWrap (N == 1 is the reference's single centre-channel run, which has no translator at all
    // -- sdrx.cpp:1408, :1684 -- and goes through the generic kernels)
    if (N <= kWrap || fast_smem_bytes(N) > 227 * 1024) return false;
    return true;
}

// Conjugate partners among the translator tables the caller supplied ([C][N] float2, row ignored where has[c] == 0):
// partner[c] = d when table[d][n] == conj(table[c][n]) for every n (compared as numbers), else -1.
inline std::vector<int> fast_find_partners(const float *tables, const uint8_t *has, int C, int N) {
    std::vector<int> partner((size_t)C, -1);
    if (!tables) return partner;
    auto bits = [](float f) { uint32_t u; memcpy(&u, &f, 4); return (u << 1) == 0 ? 0u : u; };   // +-0 alike
    auto key = [&](int c, bool conj) {
        uint64_t h = 1469598103934665603ull;
        const float *r = tables + (size_t)c * N * 2;
        for (int n = 0; n < N; ++n) {
            h = (h ^ bits(r[2 * n])) * 1099511628211ull;
            h = (h ^ bits(conj ? -r[2 * n + 1] : r[2 * n + 1])) * 1099511628211ull;
        }
        return h;
    };
    std::multimap<uint64_t, int> by_key;
    for (int c = 0; c < C; ++c) if (has[c]) by_key.insert({key(c, false), c});
    for (int c = 0; c < C; ++c) {
        if (!has[c] || partner[c] >= 0) continue;
        auto range = by_key.equal_range(key(c, true));
        for (auto it = range.first; it != range.second; ++it) {
            const int d = it->second;
            if (d == c || partner[d] >= 0) continue;
            const float *a = tables + (size_t)c * N * 2, *b = tables + (size_t)d * N * 2;
            bool ok = true;
            for (int n = 0; n < N && ok; ++n) ok = a[2 * n] == b[2 * n] && a[2 * n + 1] == -b[2 * n + 1];
            if (ok) { partner[c] = d; partner[d] = c; break; }
        }
    }
    return partner;
}

// Linearise (group, unit) work, cut it into one contiguous range per CTA, snap cuts that fall close to
// a group boundary onto it, split ranges at group boundaries into pieces, lay the two halves of every
// piece out, and list the stage-3 outputs that need tails added.
struct FastPlan {                      // the host-side result, before upload
    std::vector<FastCta> ctas;
    std::vector<FastPiece> pieces;
    std::vector<FastFix> fixes;
    std::vector<int2> fix_range;
    int max_slots = 0;
};

// One piece [ub, ub+L) of a group: halves, segment length (see fast_segment / fast_cut).
inline FastPiece fast_make_piece(int g, int ub, int L, int slot) {
    const int half_segs = kFastSegs / 2;
    FastPiece p{g, ub, ub + L, L, 0, slot};
    const int cut = ub == 0 ? kFastWarmStart - 1 : 0;            // see fast_cut
    if (L >= 2 * kFastMinSegPiece) p.H = (L - cut + 1) / 2;
    if (p.H < 1) p.H = L;
    int ns = std::max(1, std::min(half_segs, p.H / kFastMinSegPiece));
    if (cut) ns = std::max(1, std::min(ns, (p.H + cut) / (2 * cut + 1)));   // keep the segments long enough to shift the first one
    p.seg_len = std::max((p.H + cut + ns - 1) / ns, (L - p.H + ns - 1) / ns);
    if (p.seg_len <= 2 * cut) p.seg_len = std::max((p.H + ns - 1) / ns, (L - p.H + ns - 1) / ns);   // too short to shift
    return p;
}
// steps (125-sample units per thread, warm-up included) a CTA spends on a piece: every warp runs the longest segment's
inline int fast_piece_steps(const FastPiece &p) {
    return ((p.u_begin == 0 && !fast_cut(p)) ? kFastWarmStart : 1) + p.seg_len;
}

// Linearise (group, unit) work and hand every CTA a contiguous range that costs at most T steps, for the smallest
// T that needs no more CTAs than there are SMs: a range that crosses a group boundary is two pieces (the table
// in shared memory belongs to one group), each paying its own warm-up step, so ranges are sized by cost, not
// by length -- all SMs finish together.  Then lay the two halves of every piece out and list the stage-3
// outputs that need tails added.
inline int fast_plan_schedule(int groups, int U, int n_sms, FastPlan &plan) {
    const int half_segs = kFastSegs / 2;
    const long total = (long)groups * U;
    auto max_len = [&](int ub, int rem, int budget) {            // largest L <= rem whose piece fits the budget (0: none)
        int lo = 0, hi = rem;
        while (lo < hi) {
            const int mid = (lo + hi + 1) / 2;
            if (fast_piece_steps(fast_make_piece(0, ub, mid, 0)) <= budget) lo = mid; else hi = mid - 1;
        }
        return lo;
    };
    auto simulate = [&](int T, std::vector<FastCta> *ctas, std::vector<FastPiece> *pieces, int *max_slots) -> long {
        long wk = 0, n = 0;
        int slot = 0, slot_group = -1;                           // slots are numbered per group (pieces of a group are consecutive)
        while (wk < total) {
            int budget = T;
            FastCta cta{pieces ? (int)pieces->size() : 0, 0};
            bool took = false;
            while (wk < total && budget >= (took ? 3 + kFastPiecePenalty : 2)) {
                const int g = (int)(wk / U), ub = (int)(wk % U);
                // a further piece of the same CTA belongs to the next group: the table is reloaded and the staging
                // pipeline refilled, which the trace (tools/diag_bank_trace.py) puts at about one step
                const int extra = took ? kFastPiecePenalty : 0;
                int L = max_len(ub, U - ub, budget - extra);
                if (L == 0) break;
                if (took && L < kFastSegs && L < U - ub) break;                  // not worth a table reload
                if (U - ub - L > 0 && U - ub - L < 2 * kFastMinSegPiece) {       // do not leave a sliver of the group behind
                    const int all = U - ub;
                    if (fast_piece_steps(fast_make_piece(0, ub, all, 0)) <= budget - extra) L = all;
                }
                if (g != slot_group) { slot_group = g; slot = 0; }
                const FastPiece p = fast_make_piece(g, ub, L, slot);
                budget -= fast_piece_steps(p) + extra;
                slot += kFastSlotsPerPiece;
                if (pieces) { pieces->push_back(p); *max_slots = std::max(*max_slots, slot); }
                wk += L;
                took = true;
            }
            if (!took) return -1;                                // T too small to make progress
            if (ctas) { cta.piece_end = (int)pieces->size(); ctas->push_back(cta); }
            ++n;
        }
        return n;
    };
    int lo = 2, hi = kFastWarmStart + 2 + U;                     // a whole group fits one CTA at hi
    while (simulate(hi, nullptr, nullptr, nullptr) < 0 || simulate(hi, nullptr, nullptr, nullptr) > n_sms) hi *= 2;
    while (lo < hi) {
        const int mid = (lo + hi) / 2;
        const long n = simulate(mid, nullptr, nullptr, nullptr);
        if (n > 0 && n <= n_sms) hi = mid; else lo = mid + 1;
    }
    // small jobs: do not shred the block into segments that mostly warm up when there are SMs to spare anyway
    const int T = std::max(lo, std::min(1 + kFastMinSeg, (int)std::max<long>(2, 1 + (total + kFastSegs - 1) / kFastSegs)));
    std::vector<FastCta> &ctas = plan.ctas;
    std::vector<FastPiece> &pieces = plan.pieces;
    int max_slots = 0;
    if (simulate(T, &ctas, &pieces, &max_slots) < 0) return fail(SDRX_ERR_STATE, "internal: channel-bank schedule");
    // fix list: every segment end inside the block hands its four partial sums to the outputs after it
    std::vector<FastFix> &fixes = plan.fixes;
    std::vector<int2> &fix_range = plan.fix_range;
    fix_range.assign((size_t)groups, make_int2(0, 0));
    {
        std::vector<std::vector<int>> src((size_t)groups * U);
        for (const FastPiece &p : pieces)
            for (int half = 0; half < 2; ++half)
                for (int w = 0; w < half_segs; ++w) {
                    int u0, u1;
                    if (!fast_segment(p, w, half, &u0, &u1) || u1 >= U) continue;
                    for (int q = 0; q < 4 && u1 + q < U; ++q)
                        src[(size_t)p.group * U + u1 + q].push_back((p.slot_base + half * half_segs + w) * 4 + q);
                }
        for (int g = 0; g < groups; ++g) {
            fix_range[g].x = (int)fixes.size();
            for (int u = 0; u < U; ++u) {
                const std::vector<int> &v = src[(size_t)g * U + u];
                if (v.empty()) continue;
                if (v.size() > 4) return fail(SDRX_ERR_STATE, "internal: more than four tails reach one stage-3 output");
                FastFix f{g, u, (int)v.size(), {0, 0, 0, 0}};
                for (size_t k = 0; k < v.size(); ++k) f.src[k] = v[k];
                fixes.push_back(f);
            }
            fix_range[g].y = (int)fixes.size();
        }
    }
    plan.max_slots = max_slots;
    return SDRX_OK;
}

// Self-check of a plan (also exported for the CPU tests): every (group, unit) is produced by exactly one segment
// and no output receives more than four tails (a stage-2 output of unit v feeds stage-3 outputs v..v+4, so the
// segment that owns v reaches at most the four outputs after its end, whatever its length).  Returns the number
// of violations.
inline int fast_check_plan(int groups, int U, const FastPlan &plan) {
    int bad = 0;
    std::vector<int> cover((size_t)groups * U, 0);
    for (const FastPiece &p : plan.pieces)
        for (int half = 0; half < 2; ++half)
            for (int w = 0; w < kFastSegs / 2; ++w) {
                int u0, u1;
                if (!fast_segment(p, w, half, &u0, &u1)) continue;
                if (u1 <= u0 || u0 < p.u_begin || u1 > p.u_end) { ++bad; continue; }
                for (int u = u0; u < u1; ++u) ++cover[(size_t)p.group * U + u];
            }
    for (int v : cover) if (v != 1) ++bad;
    for (const FastFix &f : plan.fixes) if (f.n < 1 || f.n > 4) ++bad;
    for (const FastCta &c : plan.ctas) if (c.piece_end <= c.piece_begin) ++bad;
    return bad;
}

inline int fast_build_schedule(FastBank &fb, FastSchedule &out, int U) {
    FastPlan plan;
    int rc = fast_plan_schedule(fb.groups, U, fb.n_sms, plan);
    if (rc) return rc;
    if (fast_check_plan(fb.groups, U, plan) != 0) return fail(SDRX_ERR_STATE, "internal: channel-bank schedule does not tile the block");
    const std::vector<FastCta> &ctas = plan.ctas;
    const std::vector<FastPiece> &pieces = plan.pieces;
    const std::vector<FastFix> &fixes = plan.fixes;
    const std::vector<int2> &fix_range = plan.fix_range;
    const int max_slots = plan.max_slots;
    out.U = U;
    out.n_ctas = (int)ctas.size();
    out.n_slots = max_slots;           // tails are indexed [channel][slot][4]
    out.n_fixes = (int)fixes.size();
    CU(cudaMalloc((void **)&out.d_ctas, sizeof(FastCta) * ctas.size()));
    CU(cudaMalloc((void **)&out.d_pieces, sizeof(FastPiece) * pieces.size()));
    CU(cudaMalloc((void **)&out.d_fixes, sizeof(FastFix) * std::max<size_t>(1, fixes.size())));
    CU(cudaMemcpy(out.d_ctas, ctas.data(), sizeof(FastCta) * ctas.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(out.d_pieces, pieces.data(), sizeof(FastPiece) * pieces.size(), cudaMemcpyHostToDevice));
    if (!fixes.empty())
        CU(cudaMemcpy(out.d_fixes, fixes.data(), sizeof(FastFix) * fixes.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&out.d_fix_range, sizeof(int2) * fix_range.size()));
    CU(cudaMemcpy(out.d_fix_range, fix_range.data(), sizeof(int2) * fix_range.size(), cudaMemcpyHostToDevice));
    return SDRX_OK;
}

// Launch the bank for one block of n = 125*U samples (all stage phases zero), then -- unless the caller's next
// kernel applies the fix list itself (`deferred_fix`) -- the tail fix-up.
template <int FMT>
int fast_launch(FastBank &fb, cudaStream_t st, const void *raw, int hist, int valid_lo, int N, int pos0, unsigned n,
                float2 *s3, long s3_stride, int s3_off, uint64_t *launches, cudaEvent_t ev0 = nullptr,
                cudaEvent_t ev1 = nullptr, const FastSchedule **deferred_fix = nullptr, int tail_half = 0, long stream_bytes = 0) {
    const int U = (int)(n / kUnit);
    FastSchedule *sc = nullptr;
    for (auto &s : fb.sched) if (s.U == U) sc = &s;
    if (!sc) {
        FastSchedule &victim = fb.sched[0].U == 0 ? fb.sched[0] : fb.sched[1];
        CU(cudaDeviceSynchronize());            // a launch (or a side-stream reader) may still use the evicted schedule
        fb.free_schedule(victim);
        int rc = fast_build_schedule(fb, victim, U);
        if (rc) return rc;
        sc = &victim;
    }
    if (sc->n_slots > fb.tails_slots) {
        CU(cudaDeviceSynchronize());
        cudaFree(fb.d_tails);
        fb.d_tails = nullptr;
        CU(cudaMalloc((void **)&fb.d_tails, 2 * sizeof(float2) * 4 * (size_t)sc->n_slots * fb.C));
        fb.tails_slots = sc->n_slots;
    }
    if (ev0) CU(cudaEventRecord(ev0, st));
    if (fb.nch == 2)
        k_bank_fast<FMT, 2><<<sc->n_ctas, kFastThreads, fast_smem_bytes(N), st>>>(
            raw, hist, valid_lo, fb.d_table_slot, fb.tstride, fb.d_slot_chan, N, pos0, fb.taps, sc->d_ctas, sc->d_pieces, U,
            s3, s3_stride, s3_off, fb.tails_half(tail_half), fb.tails_slots, fb.gps, stream_bytes);
    else
        k_bank_fast<FMT, 1><<<sc->n_ctas, kFastThreads, fast_smem_bytes(N), st>>>(
            raw, hist, valid_lo, fb.d_table_slot, fb.tstride, fb.d_slot_chan, N, pos0, fb.taps, sc->d_ctas, sc->d_pieces, U,
            s3, s3_stride, s3_off, fb.tails_half(tail_half), fb.tails_slots, fb.gps, stream_bytes);
    if (ev1) CU(cudaEventRecord(ev1, st));
    ++*launches;
    if (deferred_fix) *deferred_fix = sc;           // the caller's next kernel applies the fix list itself
    else if (sc->n_fixes > 0) {
        const int per_group = kFastCh * fb.nch;
        const int threads = sc->n_fixes * per_group;
        k_fix_tails<<<div_up(threads, 256), 256, 0, st>>>(s3, s3_stride, s3_off, fb.tails_half(tail_half), fb.tails_slots, sc->d_fixes,
                                                          sc->n_fixes, fb.d_slot_chan, per_group);
        ++*launches;
    }
    CU(cudaGetLastError());
    return SDRX_OK;
}

// tables: the caller's translator tables [C][N] float2 (host) or nullptr; has[c] == 0 -> no translator (the slot's
// table column is (1, 0): multiplying by it is exact).
// C channels PER STREAM; every stream carries the same channel plan (tables), channel c of stream s is engine channel s * C + c.
inline int fast_init(FastBank &fb, int C, int N, const float *h1, const float *h2, const float *h3, const float *tables,
                     const uint8_t *has, bool allow_pairs, int n_streams = 1) {
    fb.C = C * n_streams;
    fb.n_streams = n_streams;
    fb.stages_fused = 3;
    {
        int a = N, b = kUnit;
        while (b) { const int t = a % b; a = b; b = t; }
        fb.period_units = N / a;
    }
    for (int i = 0; i < 17; ++i) fb.taps.h1[i] = make_float2(h1[i], h1[i]);
    for (int i = 0; i < 19; ++i) fb.taps.h2[i] = make_float2(h2[i], h2[i]);
    for (int i = 0; i < 24; ++i) fb.taps.h3[i] = i < 21 ? make_float2(h3[i], h3[i]) : make_float2(0.0f, 0.0f);
    int dev = 0;
    CU(cudaGetDevice(&dev));
    CU(cudaDeviceGetAttribute(&fb.n_sms, cudaDevAttrMultiProcessorCount, dev));
    fb.n_sms = std::max(fb.n_sms - kFastSideSms, 1);        // SMs left to the latency-bound side kernels (demod.cuh kSideSms)

    // ---- slots: conjugate pairs share a slot when enough channels pair up to shrink the work
    std::vector<uint8_t> has_v((size_t)C, 0);
    for (int c = 0; c < C; ++c) has_v[c] = (tables && (has ? has[c] != 0 : true)) ? 1 : 0;
    std::vector<int> partner = allow_pairs ? fast_find_partners(tables, has_v.data(), C, N) : std::vector<int>((size_t)C, -1);
    int n_pairs = 0;
    for (int c = 0; c < C; ++c) if (partner[c] > c) ++n_pairs;
    fb.nch = (n_pairs * 4 >= C) ? 2 : 1;            // pairing costs a second accumulator set for every slot
    fb.n_pairs = fb.nch == 2 ? n_pairs : 0;
    std::vector<int> slot_chan;
    for (int c = 0; c < C; ++c) {
        if (fb.nch == 2) {
            if (partner[c] >= 0 && partner[c] < c) continue;          // rides in its partner's slot
            slot_chan.push_back(c);
            slot_chan.push_back(partner[c]);
        } else {
            slot_chan.push_back(c);
        }
    }
    const int slots = (int)slot_chan.size() / fb.nch;
    fb.gps = (slots + kFastCh - 1) / kFastCh;
    fb.groups = fb.gps * n_streams;
    slot_chan.resize((size_t)fb.gps * kFastCh * fb.nch, -1);
    fb.tstride = fb.gps * kFastCh;
    std::vector<float2> tslot((size_t)N * fb.tstride, make_float2(1.0f, 0.0f));
    std::vector<int> chan_group((size_t)C * n_streams, 0);
    for (int s = 0; s < fb.gps * kFastCh; ++s) {
        for (int q = 0; q < fb.nch; ++q) {
            const int c = slot_chan[(size_t)s * fb.nch + q];
            if (c >= 0)
                for (int k = 0; k < n_streams; ++k) chan_group[(size_t)k * C + c] = k * fb.gps + s / kFastCh;
        }
        const int c0 = slot_chan[(size_t)s * fb.nch];
        if (c0 < 0 || !has_v[c0]) continue;
        const float *r = tables + (size_t)c0 * N * 2;
        for (int n = 0; n < N; ++n) tslot[(size_t)n * fb.tstride + s] = make_float2(r[2 * n], r[2 * n + 1]);
    }
    CU(cudaMalloc((void **)&fb.d_table_slot, sizeof(float2) * tslot.size()));
    CU(cudaMemcpy(fb.d_table_slot, tslot.data(), sizeof(float2) * tslot.size(), cudaMemcpyHostToDevice));
    {   // the slots of stream k carry the same channels, k * C further
        const size_t one = slot_chan.size();
        slot_chan.resize(one * n_streams);
        for (int k = 1; k < n_streams; ++k)
            for (size_t i = 0; i < one; ++i) slot_chan[k * one + i] = slot_chan[i] < 0 ? -1 : slot_chan[i] + k * C;
    }
    CU(cudaMalloc((void **)&fb.d_slot_chan, sizeof(int) * slot_chan.size()));
    CU(cudaMemcpy(fb.d_slot_chan, slot_chan.data(), sizeof(int) * slot_chan.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&fb.d_chan_group, sizeof(int) * chan_group.size()));
    CU(cudaMemcpy(fb.d_chan_group, chan_group.data(), sizeof(int) * chan_group.size(), cudaMemcpyHostToDevice));

    {
        const size_t smem = fast_smem_bytes(N);
        const void *fn[6] = {(const void *)k_bank_fast<SDRX_FMT_U8, 1>, (const void *)k_bank_fast<SDRX_FMT_S16, 1>,
                             (const void *)k_bank_fast<SDRX_FMT_F32, 1>, (const void *)k_bank_fast<SDRX_FMT_U8, 2>,
                             (const void *)k_bank_fast<SDRX_FMT_S16, 2>, (const void *)k_bank_fast<SDRX_FMT_F32, 2>};
        for (const void *f : fn) { int rc = raise_dyn_smem(f, smem); if (rc) return rc; }
    }
    fb.enabled = true;
    return SDRX_OK;
}

}  // namespace sdrx

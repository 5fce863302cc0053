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


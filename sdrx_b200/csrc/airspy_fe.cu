// sdrx_b200 -- Airspy real -> IQ front end on the GPU: the step libairspy runs on the host
// in front of the sdrx hot path for AIRSPY_SAMPLE_FLOAT32_IQ (airspy.c:393-396).  Moving it here lets the wire format
// be the ADC's own 12-bit samples (2 B, or 1.5 B packed, per REAL sample at 20 MS/s = 3-4 B per IQ sample instead
// of 8) and hands the channel bank a device-resident float IQ block.
//
// Restated (paths under libairspy/libairspy/src of the reference):
//   airspy.c:323-338  unpack_samples          -> load_adc8<true>
//   airspy.c:311-321  convert_samples_float   -> (raw - 2048) * (1/2048)
//   iqconverter_float.c:455-468 remove_dc     -> k_fe_dc: s -= avg; avg += 0.01f * s over ALL samples in order
//   iqconverter_float.c:470-502 translate_fs_4, :191-230 fir_interleaved_24, :430-451 delay_interleaved -> k_fe_iq
//
// remove_dc is one sequential recurrence over the whole stream: avg' = avg + 0.01f * (x - avg), i.e. the affine map
// avg' = a * avg + 0.01f * x with a = 1 - 0.01f.  The block is cut into 128-sample segments, one thread each, and
// the recurrence is run in three short passes instead of one long one:
//   k_fe_partial  every segment from avg = 0: B[k], so that the segment maps avg_in -> a^128 * avg_in + B[k];
//   k_fe_scan     one small CTA composes those maps into the avg at every segment start, starting
//                 from the carried block-start avg -- right to a few ulp, not yet the reference's rounded sequence;
//   k_fe_dc       every thread starts kFeBack = 3 segments early from that estimate and runs the reference's own
//                 float recurrence: it is a contraction (0.99^384 = 0.02), so the few-ulp error dies out and the
//                 trajectory snaps onto the reference's rounded sequence before the thread's own segment begins
//                 (segments 0..3 start from the exact carried avg).
// A thread's dependent chain is 128 + 512 steps instead of 640 000.  SDRX_FE_EXACT_DC runs the recurrence on one
// thread instead (tests): bit-exact by construction.
// Everything after remove_dc is a pure function of a short window and is evaluated in the reference's order,
// un-contracted: bit-exact given bit-exact inputs.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"
#include "host_common.h"

using namespace sdrx;

namespace {

constexpr int kFeSeg = 128;          // real samples per thread of k_fe_dc
constexpr int kFeBack = 3;           // warm-up segments in front of a thread's own
constexpr int kFeThreads = 128;      // CTA size of the block-wide passes
constexpr int kFeMaxCtas = 1 << 20;  // ... and their largest grid (the passes walk the block with a grid stride): see fe_run
constexpr int kFeEvenHist = 24;      // >= len - 1 = 23 rotated even samples carried between blocks
constexpr int kFeOddHist = 12;       // len / 2 delayed odd samples
constexpr int kFeMaxLen = 64;

// Debug trace (tools/diag_fe_chain.py): when set, kernel k records the globaltimer at which its CTA 0 started in
// slot 2k and the latest CTA end in slot 2k + 1 (k: 0 partial, 1 scan, 2 dc, 3 iq, 4 carry).
__device__ unsigned long long *g_fe_trace = nullptr;
__device__ __forceinline__ unsigned long long fe_now() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void fe_mark_start(int k) {
    if (g_fe_trace && threadIdx.x == 0 && blockIdx.x == 0) g_fe_trace[2 * k] = fe_now();
}
__device__ __forceinline__ void fe_mark_end(int k) {
    if (g_fe_trace && threadIdx.x == 0) atomicMax(&g_fe_trace[2 * k + 1], fe_now());
}

// eight consecutive ADC samples starting at index `idx` (a multiple of 8) of a buffer in wire format
template <bool PACKED>
__device__ __forceinline__ void load_adc8(const void *__restrict__ raw, long idx, int (&v)[8]) {
    if constexpr (PACKED) {                      // airspy.c:323-338
        const uint32_t *in = reinterpret_cast<const uint32_t *>(raw) + (idx >> 3) * 3;
        const uint32_t a = __ldg(in), b = __ldg(in + 1), c = __ldg(in + 2);
        v[0] = (a >> 20) & 0xfff;
        v[1] = (a >> 8) & 0xfff;
        v[2] = ((a & 0xff) << 4) | ((b >> 28) & 0xf);
        v[3] = (b >> 16) & 0xfff;
        v[4] = (b >> 4) & 0xfff;
        v[5] = ((b & 0xf) << 8) | ((c >> 24) & 0xff);
        v[6] = (c >> 12) & 0xfff;
        v[7] = c & 0xfff;
    } else {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const uint16_t *>(raw) + idx));
        v[0] = q.x & 0xffff; v[1] = q.x >> 16; v[2] = q.y & 0xffff; v[3] = q.y >> 16;
        v[4] = q.z & 0xffff; v[5] = q.z >> 16; v[6] = q.w & 0xffff; v[7] = q.w >> 16;
    }
}

__device__ __forceinline__ float adc_to_float(int v) {            // airspy.c:311-321, SAMPLE_SCALE = 1/2048
    return __fmul_rn((float)(v - 2048), 1.0f / 2048.0f);
}

// One remove_dc step (iqconverter_float.c:461-464).
__device__ __forceinline__ float dc_step(float x, float &avg) {
    const float s = __fsub_rn(x, avg);
    avg = __fadd_rn(avg, __fmul_rn(0.01f, s));
    return s;
}

// remove_dc over samples [w0, s1) of the block starting from `avg`; with EMIT, the samples from s0 on also go through
// translate_fs_4's sign / hbc pattern into ev / od.  Groups of eight samples travel through a ring of registers,
// fetched kFeAhead groups (64 samples, ~770 cycles of the recurrence) ahead of their use: the loads do not depend on
// the recurrence, so its dependent chain never waits for memory.  A ring slot holds the eight samples as four words
// of two uint16 each, whatever the wire format.
template <bool PACKED, bool EMIT>
__device__ __forceinline__ float fe_walk(const void *__restrict__ raw, long w0, long s0, long s1, float avg, float hbc,
                                         float *__restrict__ ev, float *__restrict__ od) {
    constexpr int kFeAhead = 8;
    auto fetch = [&](long i) -> uint4 {
        if (i >= s1) return make_uint4(0, 0, 0, 0);
        if constexpr (PACKED) {
            int v[8];
            load_adc8<true>(raw, i, v);
            return make_uint4(v[0] | (v[1] << 16), v[2] | (v[3] << 16), v[4] | (v[5] << 16), v[6] | (v[7] << 16));
        } else {
            return __ldg(reinterpret_cast<const uint4 *>(reinterpret_cast<const uint16_t *>(raw) + i));
        }
    };
    uint4 ring[kFeAhead];
#pragma unroll
    for (int j = 0; j < kFeAhead; ++j) ring[j] = fetch(w0 + 8L * j);
    const long ng = (s1 - w0) >> 3;
    for (long g0 = 0; g0 < ng; g0 += kFeAhead) {
#pragma unroll
        for (int j = 0; j < kFeAhead; ++j) {
            const long i = w0 + 8 * (g0 + j);
            if (g0 + j < ng) {
                const uint4 q = ring[j];
                ring[j] = fetch(i + 8L * kFeAhead);
                const int v[8] = {(int)(q.x & 0xffff), (int)(q.x >> 16), (int)(q.y & 0xffff), (int)(q.y >> 16),
                                  (int)(q.z & 0xffff), (int)(q.z >> 16), (int)(q.w & 0xffff), (int)(q.w >> 16)};
                // (raw - 2048) * (1/2048) without the integer->float converter on the way: 2^23 + v is the float whose
                // low mantissa bits are v; subtracting 2^23 + 2048 and scaling by 2^-11 are both exact, so the value is
                // bit for bit convert_samples_float's.  All eight conversions sit in front of the recurrence's chain.
                float x[8], s[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    x[k] = __fmul_rn(__fsub_rn(__int_as_float(0x4B000000 | v[k]), 8390656.0f), 1.0f / 2048.0f);
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] = dc_step(x[k], avg);
                if (EMIT && i >= s0) {
                    // translate_fs_4 (iqconverter_float.c:489-498): [-s0, -s1*hbc, s2, s3*hbc] per group of four
                    const long p = i >> 1;
                    float4 e, o;
                    e.x = -s[0];               o.x = __fmul_rn(-s[1], hbc);
                    e.y = s[2];                o.y = __fmul_rn(s[3], hbc);
                    e.z = -s[4];               o.z = __fmul_rn(-s[5], hbc);
                    e.w = s[6];                o.w = __fmul_rn(s[7], hbc);
                    *reinterpret_cast<float4 *>(ev + kFeEvenHist + p) = e;
                    *reinterpret_cast<float4 *>(od + kFeOddHist + p) = o;
                }
            }
        }
    }
    return avg;
}

// Pass 1: the recurrence over one segment from avg = 0 (the additive term of the segment's affine map).
template <bool PACKED>
static __global__ void __launch_bounds__(kFeThreads)
k_fe_partial(const void *__restrict__ raw, int n, float *__restrict__ seg_b) {
    fe_mark_start(0);
    const int whole = n / kFeSeg;                   // only whole segments feed the scan (the last start needs no successor)
    for (int seg = blockIdx.x * blockDim.x + threadIdx.x; seg < whole; seg += gridDim.x * blockDim.x) {
        const long s0 = (long)seg * kFeSeg;
        seg_b[seg] = fe_walk<PACKED, false>(raw, s0, s0, s0 + kFeSeg, 0.0f, 0.0f, nullptr, nullptr);
    }
    fe_mark_end(0);
}

// Pass 2: avg at every segment start, start[0] = the carried avg, start[k+1] = A * start[k] + B[k] (A = a^kFeSeg),
// composed by one CTA of 128 threads: each thread folds a contiguous chunk, thread 0 chains the 128 chunks, every
// thread re-walks its chunk from its chunk start.  Single precision: the estimate only has to be right to a few ulp.
static __global__ void __launch_bounds__(128)
k_fe_scan(const float *__restrict__ seg_b, int nseg, const float *__restrict__ avg_io, float A, float *__restrict__ seg_start) {
    __shared__ float sa[128], sb[128], st[128];
    fe_mark_start(1);
    const int lane = threadIdx.x;
    const int m = (nseg + 127) / 128;
    const int k0 = lane * m, k1 = min(nseg, k0 + m);
    float a = 1.0f, b = 0.0f;
    for (int k = k0; k < k1; ++k) { b = fmaf(A, b, seg_b[k]); a *= A; }
    sa[lane] = a; sb[lane] = b;
    __syncthreads();
    if (lane == 0) {
        float v = avg_io[0];
        for (int l = 0; l < 128; ++l) { st[l] = v; v = fmaf(sa[l], v, sb[l]); }
    }
    __syncthreads();
    float v = st[lane];
    for (int k = k0; k < k1; ++k) { seg_start[k] = v; v = fmaf(A, v, seg_b[k]); }
    fe_mark_end(1);
}

// Pass 3: remove_dc + translate_fs_4's sign / hbc pattern, segment-parallel.
//   raw        this block, wire format, n samples (n % 8 == 0)
//   seg_start  estimate of avg at every segment start (k_fe_scan)
//   avg_io     [0] avg at the block start (in), [1] avg at the block end (out)
//   ev / od    rotated even / odd samples of the block, behind their carried history prefixes
template <bool PACKED>
static __global__ void __launch_bounds__(kFeThreads)
k_fe_dc(const void *__restrict__ raw, const float *__restrict__ seg_start, int n, float *__restrict__ avg_io,
        float hbc, float *__restrict__ ev, float *__restrict__ od, int exact) {
    const int gt = blockIdx.x * blockDim.x + threadIdx.x;
    fe_mark_start(2);
    if (exact) {
        if (gt == 0) avg_io[1] = fe_walk<PACKED, true>(raw, 0, 0, n, avg_io[0], hbc, ev, od);
        return;
    }
    for (int seg = gt; (long)seg * kFeSeg < n; seg += gridDim.x * blockDim.x) {
        const long s0 = (long)seg * kFeSeg;
        const long s1 = s0 + kFeSeg < n ? s0 + kFeSeg : n;
        const int kb = seg <= kFeBack ? 0 : seg - kFeBack;              // segment the warm-up starts at
        const float avg0 = kb == 0 ? avg_io[0] : seg_start[kb];         // from the block start the trajectory is the exact one
        const float avg = fe_walk<PACKED, true>(raw, (long)kb * kFeSeg, s0, s1, avg0, hbc, ev, od);
        if (s1 == n) avg_io[1] = avg;
    }
    fe_mark_end(2);
}

// ---- the same two passes in their NARROW form (SDRX_FE_BESIDE_ENGINE): few threads, long chains, input staged ----
// A front end that feeds a channel bank on the same GPU has to live in the 3 SMs the bank's 145-CTA launch leaves free
// (profiles/r01y_frontend_timeline.txt).  The wide passes above are ~600 SM-us of work per block; this form cuts the
// recurrences to 3x the stream (512-sample segments, one warm-up segment: 0.99^512 = 0.006) on 1250 threads, and
// gets the segments -- which lie 1 KB apart, so a thread reading its own would touch a cache line per lane -- through
// shared memory: the CTA copies the next 32 samples of all its segments with coalesced 16 B cp.async (8 segments
// per warp instruction, whole sectors), three tiles deep, rows padded to 80 B (conflict-free 128-bit reads).  The
// outputs go back the same way: a warp parks its 32 rows of ev / od in shared memory and writes them out 64 B per
// segment (8 segments per store instruction instead of 32 half sectors).
// CTA size, measured on the 960-channel chain (tools/bench_frontend.py): 512 threads (3 CTAs, 205 KB each: every one
// needs an SM to itself) 197 us per block, 256 threads 180, 128 threads 172, 64 threads 171 -- small CTAs slip in
// next to the engine's side-stream kernels and, because the consumer wait only guards the last pass, mostly run
// ahead of the bank launch they feed.  128 threads: 10 CTAs of 51 KB.
constexpr int kNSeg = 512, kNBack = 1, kNThreads = 128, kNTile = 32, kNStages = 3;
constexpr int kNRow = 80;                                      // bytes per staged row: 64 (u16) or 48 (packed) + pad
constexpr int kNORow = 80;                                     // ... and per output row (64 B of ev or od + pad)
inline size_t fe_narrow_smem() { return (size_t)kNStages * kNThreads * kNRow + (size_t)kNThreads * 2 * kNORow; }

// EMIT = false: pass 1 (seg_b); EMIT = true: pass 3 (ev / od, avg_io[1]).  Tiles are counted from base = s0 - 512 * EMIT.
// Sample and byte indices are 32-bit here (sdrx_airspy_fe_create bounds max_samples for this form); everything that
// does not change from tile to tile is worked out once, in front of the loop.
template <bool PACKED, bool EMIT>
static __global__ void __launch_bounds__(kNThreads)
k_fe_walk_narrow(const void *__restrict__ raw, int n, const float *__restrict__ seg_start, float *__restrict__ avg_io,
                 float hbc, float *__restrict__ ev, float *__restrict__ od, float *__restrict__ seg_b) {
    extern __shared__ __align__(16) unsigned char stage[];
    unsigned char *const obuf = stage + (size_t)kNStages * kNThreads * kNRow;     // [thread][ev, od][kNORow]
    fe_mark_start(EMIT ? 2 : 0);
    constexpr int kBytes = PACKED ? kNTile * 3 / 2 : kNTile * 2;       // bytes of one row of a tile
    constexpr int kChunks = kBytes / 16;
    constexpr int kTiles = (EMIT ? kNSeg * (1 + kNBack) : kNSeg) / kNTile;
    constexpr int kGroups = kNTile / 8;
    const int t = threadIdx.x;
    const int seg0 = blockIdx.x * kNThreads;
    const int nseg = EMIT ? (n + kNSeg - 1) / kNSeg : n / kNSeg;
    const int back = EMIT ? kNSeg * kNBack : 0;
    const uint8_t *rb = reinterpret_cast<const uint8_t *>(raw);
    const int end_byte = PACKED ? n / 2 * 3 : n * 2;

    // loader: chunk `it` of this thread is (row, ch) = ((t + it * kNThreads) / kChunks, ... % kChunks) in every tile
    int g_ofs[kChunks];                        // byte offset of the chunk in tile 0 (negative in front of the block)
    unsigned s_ofs[kChunks];                   // its place inside a stage
    bool l_ok[kChunks];
#pragma unroll
    for (int it = 0; it < kChunks; ++it) {
        const int q = t + it * kNThreads, row = q / kChunks, ch = q % kChunks;
        const int smp = (seg0 + row) * kNSeg - back;
        g_ofs[it] = (PACKED ? smp / 2 * 3 : smp * 2) + ch * 16;
        s_ofs[it] = (unsigned)(row * kNRow + ch * 16);
        l_ok[it] = seg0 + row < nseg;
    }
    const unsigned stage_s = (unsigned)__cvta_generic_to_shared(stage);
    auto issue = [&](int j) {
        if (j < kTiles) {
            const unsigned slot = (unsigned)(j % kNStages) * kNThreads * kNRow;
#pragma unroll
            for (int it = 0; it < kChunks; ++it) {
                const int byte = g_ofs[it] + j * kBytes;
                if (!l_ok[it] || byte < 0 || byte >= end_byte) continue;
                if (byte + 16 <= end_byte) {
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stage_s + slot + s_ofs[it]), "l"(rb + byte) : "memory");
                } else {                                            // packed words: the buffer may end inside a chunk
                    for (int b = 0; byte + b < end_byte; b += 4)
                        *reinterpret_cast<uint32_t *>(stage + slot + s_ofs[it] + b) = *reinterpret_cast<const uint32_t *>(rb + byte + b);
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    // walker: this thread's segment
    const int seg = seg0 + t;
    const bool mine = seg < nseg;
    const int s0 = seg * kNSeg;
    const int s1 = EMIT ? min(s0 + kNSeg, n) : s0 + kNSeg;
    const int base = s0 - back;
    int lo = s0;                                                // first sample this thread walks
    float avg = 0.0f;
    if (EMIT && mine) {
        const int kb = seg <= kNBack ? 0 : seg - kNBack;
        lo = kb * kNSeg;
        avg = kb == 0 ? avg_io[0] : seg_start[kb];
    }
    // write-out: lane (r, g) of pass k copies group g of row wbase + 8k + lane / 4 (kGroups == 4)
    static_assert(kGroups == 4, "the write-out below maps four lanes to a row");
    const int lane = t & 31, wbase = t & ~31, og = lane & 3;
    int o_i0[kGroups], o_end[kGroups];         // first sample index of the group in tile 0 / end of the row's segment
#pragma unroll
    for (int k = 0; k < kGroups; ++k) {
        const int r = wbase + 8 * k + (lane >> 2);
        const int rs0 = (seg0 + r) * kNSeg;
        o_i0[k] = rs0 - back + 8 * og;
        o_end[k] = seg0 + r < nseg ? min(rs0 + kNSeg, n) : 0;
    }

    static_assert(kNStages == 3, "the wait below counts two tiles in flight");
    issue(0); issue(1);
    for (int j = 0; j < kTiles; ++j) {
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        const unsigned char *row = stage + (size_t)(j % kNStages) * kNThreads * kNRow + t * kNRow;
        if (mine && base + (j + 1) * kNTile > lo && base + j * kNTile < s1) {
            uint32_t w[kBytes / 4];                               // the row of this tile, read with conflict-free 128-bit loads
#pragma unroll
            for (int c = 0; c < kChunks; ++c) {
                const uint4 q = *reinterpret_cast<const uint4 *>(row + 16 * c);
                w[4 * c] = q.x; w[4 * c + 1] = q.y; w[4 * c + 2] = q.z; w[4 * c + 3] = q.w;
            }
#pragma unroll
            for (int g = 0; g < kGroups; ++g) {
                const int i = base + j * kNTile + 8 * g;
                if (i < lo || i >= s1) continue;
                int v[8];
                if constexpr (PACKED) {
                    const uint32_t a = w[3 * g], b = w[3 * g + 1], c = w[3 * g + 2];
                    v[0] = (a >> 20) & 0xfff; v[1] = (a >> 8) & 0xfff; v[2] = ((a & 0xff) << 4) | ((b >> 28) & 0xf);
                    v[3] = (b >> 16) & 0xfff; v[4] = (b >> 4) & 0xfff; v[5] = ((b & 0xf) << 8) | ((c >> 24) & 0xff);
                    v[6] = (c >> 12) & 0xfff; v[7] = c & 0xfff;
                } else {
                    const uint32_t a = w[4 * g], b = w[4 * g + 1], c = w[4 * g + 2], d = w[4 * g + 3];
                    v[0] = a & 0xffff; v[1] = a >> 16; v[2] = b & 0xffff; v[3] = b >> 16;
                    v[4] = c & 0xffff; v[5] = c >> 16; v[6] = d & 0xffff; v[7] = d >> 16;
                }
                // (raw - 2048) / 2048 in one exact step: (2^23 + v) * 2^-11 - 4097 == v / 2048 - 1, representable, so the
                // fused multiply-add rounds nothing and the value is convert_samples_float's bit for bit
                float x[8], sv[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) x[k] = __fmaf_rn(__int_as_float(0x4B000000 | v[k]), 1.0f / 2048.0f, -4097.0f);
#pragma unroll
                for (int k = 0; k < 8; ++k) sv[k] = dc_step(x[k], avg);
                if (EMIT && i >= s0) {
                    float4 e, o;
                    e.x = -sv[0];              o.x = __fmul_rn(-sv[1], hbc);
                    e.y = sv[2];               o.y = __fmul_rn(sv[3], hbc);
                    e.z = -sv[4];              o.z = __fmul_rn(-sv[5], hbc);
                    e.w = sv[6];               o.w = __fmul_rn(sv[7], hbc);
                    *reinterpret_cast<float4 *>(obuf + (2 * t) * kNORow + 16 * g) = e;
                    *reinterpret_cast<float4 *>(obuf + (2 * t + 1) * kNORow + 16 * g) = o;
                }
            }
        }
        if (EMIT && j * kNTile >= back) {                        // this tile lies in the segments' own range: write it out
            __syncwarp();
#pragma unroll
            for (int k = 0; k < kGroups; ++k) {
                const int i = o_i0[k] + j * kNTile;
                if (i < o_end[k]) {
                    const int r = wbase + 8 * k + (lane >> 2);
                    *reinterpret_cast<float4 *>(ev + kFeEvenHist + (i >> 1)) = *reinterpret_cast<const float4 *>(obuf + (2 * r) * kNORow + 16 * og);
                    *reinterpret_cast<float4 *>(od + kFeOddHist + (i >> 1)) = *reinterpret_cast<const float4 *>(obuf + (2 * r + 1) * kNORow + 16 * og);
                }
            }
            __syncwarp();
        }
        issue(j + kNStages - 1);
    }
    if (mine) {
        if (!EMIT) seg_b[seg] = avg;
        else if (s1 == n) avg_io[1] = avg;
    }
    fe_mark_end(EMIT ? 2 : 0);
}

// Half-band FIR on the even samples (fir_interleaved_24's order: k[0]*(q0+q23) + k[1]*(q1+q22) + ... left to right),
// the odd samples through their 12-deep delay line.  One thread per EIGHT consecutive IQ pairs: their windows overlap
// in all but one sample each, so 32 even samples are read once (eight 16 B loads) instead of 192 separate ones.
// Specialised for libairspy's kernel (len == 24 == kFeEvenHist: the window of pair p is ev[p .. p + 23] + the newest).
static __global__ void __launch_bounds__(kFeThreads)
k_fe_iq(const float *__restrict__ ev, const float *__restrict__ od, int npair, const float *__restrict__ k, int len,
        float2 *__restrict__ out) {
  fe_mark_start(3);
  for (int p0 = (blockIdx.x * blockDim.x + threadIdx.x) * 8; p0 < npair; p0 += gridDim.x * blockDim.x * 8) {
    if (len == 24 && p0 + 8 <= npair) {
        float w[32];                                     // w[i] = ev[p0 + i]: the newest sample of pair p0 + j is w[24 + j]
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 v = *reinterpret_cast<const float4 *>(ev + p0 + 4 * i);
            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
        }
        float kk[12];
#pragma unroll
        for (int i = 0; i < 12; ++i) kk[i] = __ldg(k + i);
        const float4 o0 = *reinterpret_cast<const float4 *>(od + p0), o1 = *reinterpret_cast<const float4 *>(od + p0 + 4);
        const float oq[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
        float2 y[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float acc = __fmul_rn(kk[0], __fadd_rn(w[24 + j], w[1 + j]));
#pragma unroll
            for (int i = 1; i < 12; ++i) acc = __fadd_rn(acc, __fmul_rn(kk[i], __fadd_rn(w[24 + j - i], w[1 + j + i])));
            y[j] = make_float2(acc, oq[j]);              // od[kFeOddHist + p - 12] == od[p]
        }
        float4 *o4 = reinterpret_cast<float4 *>(out + p0);
#pragma unroll
        for (int j = 0; j < 8; j += 2) o4[j >> 1] = make_float4(y[j].x, y[j].y, y[j + 1].x, y[j + 1].y);
        continue;
    }
    const int half = len >> 1;
    for (int p = p0; p < npair && p < p0 + 8; ++p) {
        const float *q = ev + kFeEvenHist + p;
        float acc = __fmul_rn(__ldg(k), __fadd_rn(q[0], q[-(len - 1)]));
        for (int i = 1; i < half; ++i) acc = __fadd_rn(acc, __fmul_rn(__ldg(k + i), __fadd_rn(q[-i], q[-(len - 1 - i)])));
        out[p] = make_float2(acc, od[kFeOddHist + p - half]);
    }
  }
  fe_mark_end(3);
}

// Block-to-block carries: the even / odd history prefixes and avg.
static __global__ void __launch_bounds__(64)
k_fe_carry(int n, float *__restrict__ ev, float *__restrict__ od, float *__restrict__ avg_io) {
    // the even / odd prefixes slide in place: every read below happens before the barrier, every write after
    // (one CTA does it: kFeEvenHist + kFeOddHist <= 256)
    fe_mark_start(4);
    if (blockIdx.x == 0) {
        const int npair = n >> 1;
        float keep = 0.0f;
        const int t = threadIdx.x;
        if (t < kFeEvenHist) keep = ev[npair + t];
        else if (t < kFeEvenHist + kFeOddHist) keep = od[npair + (t - kFeEvenHist)];
        __syncthreads();
        if (t < kFeEvenHist) ev[t] = keep;
        else if (t < kFeEvenHist + kFeOddHist) od[t - kFeEvenHist] = keep;
        if (t == 0) avg_io[0] = avg_io[1];
    }
    fe_mark_end(4);
}

}  // namespace

struct sdrx_airspy_fe {
    int dev = 0;
    unsigned flags = 0;
    unsigned cap = 0;                    // real samples per call at most
    int len = 0;                         // hb_len / 2 + 1
    float hbc = 0.0f;
    cudaStream_t st = nullptr;
    cudaEvent_t ev_done = nullptr;
    uint8_t *d_raw = nullptr;            // wire-format staging of one call
    uint8_t *h_stage = nullptr;          // pinned staging for pageable callers
    float *d_seg_b = nullptr, *d_seg_start = nullptr;   // per-segment affine terms / avg estimates (k_fe_partial, k_fe_scan)
    double A = 0.0;                      // (1 - 0.01f)^kFeSeg (or ^kNSeg: narrow)
    bool narrow = false;                 // SDRX_FE_BESIDE_ENGINE
    cudaEvent_t out_free = nullptr;      // sdrx_airspy_fe_wait_event: the next call's output write waits for this
    float *d_ev = nullptr, *d_od = nullptr, *d_k = nullptr, *d_avg = nullptr;
    float2 *d_iq[2] = {nullptr, nullptr};    // ring: the consumer of call k may still read while call k+1 runs
    uint64_t total = 0;                  // real samples seen so far
    uint64_t calls = 0;
    unsigned last_pairs = 0;
    uint64_t launches = 0;
};

extern "C" void sdrx_airspy_fe_destroy(sdrx_airspy_fe *f) {
    if (!f) return;
    DevGuard g(f->dev);
    if (f->st) cudaStreamSynchronize(f->st);
    cudaFree(f->d_raw); cudaFree(f->d_seg_b); cudaFree(f->d_seg_start); cudaFree(f->d_ev); cudaFree(f->d_od);
    cudaFree(f->d_k); cudaFree(f->d_avg); cudaFree(f->d_iq[0]); cudaFree(f->d_iq[1]);
    if (f->h_stage) cudaFreeHost(f->h_stage);
    if (f->ev_done) cudaEventDestroy(f->ev_done);
    if (f->st) cudaStreamDestroy(f->st);
    delete f;
}

extern "C" int sdrx_airspy_fe_reset(sdrx_airspy_fe *f) {
    if (!f) return fail(SDRX_ERR_ARG, "null front end");
    DevGuard g(f->dev);
    CU(cudaStreamSynchronize(f->st));
    CU(cudaMemsetAsync(f->d_ev, 0, sizeof(float) * (kFeEvenHist + f->cap / 2 + 8), f->st));     // iqconverter_float_reset
    CU(cudaMemsetAsync(f->d_od, 0, sizeof(float) * (kFeOddHist + f->cap / 2 + 8), f->st));
    CU(cudaMemsetAsync(f->d_avg, 0, sizeof(float) * 2, f->st));
    CU(cudaStreamSynchronize(f->st));
    f->total = 0;
    f->calls = 0;
    f->last_pairs = 0;
    f->out_free = nullptr;
    return SDRX_OK;
}

extern "C" int sdrx_airspy_fe_create(int device, const float *hb_kernel, int hb_len, unsigned max_samples, unsigned flags,
                                     sdrx_airspy_fe **out) {
    if (!out || !hb_kernel) return fail(SDRX_ERR_ARG, "null argument");
    *out = nullptr;
    if (hb_len < 3 || (hb_len & 1) == 0) return fail(SDRX_ERR_ARG, "half-band kernel length must be odd");
    const int len = hb_len / 2 + 1;
    if ((len & 1) || len - 1 > kFeEvenHist || len / 2 > kFeOddHist || len > kFeMaxLen)
        return fail(SDRX_ERR_UNSUPPORTED, "half-band kernels other than libairspy's 47-tap family (len/2+1 even, <= 24) are not implemented");
    if (max_samples < 8 || max_samples % 8) return fail(SDRX_ERR_ARG, "max_samples must be a positive multiple of 8");
    if ((flags & SDRX_FE_BESIDE_ENGINE) && max_samples > (1u << 28))
        return fail(SDRX_ERR_ARG, "SDRX_FE_BESIDE_ENGINE: max_samples above 2^28 (32-bit indices in the narrow passes)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        cudaGetLastError();
        return fail(SDRX_ERR_CUDA, "no CUDA device: sdrx_b200 has no CPU path");
    }
    if (device < 0 || device >= ndev) return fail(SDRX_ERR_ARG, "device ordinal out of range");
    DevGuard g(device);
    if (!g.ok) return fail(SDRX_ERR_CUDA, "cudaSetDevice failed");
    sdrx_airspy_fe *f = new sdrx_airspy_fe();
    f->dev = device; f->flags = flags; f->cap = max_samples; f->len = len;

    f->hbc = hb_kernel[hb_len / 2];                              // iqconverter_float.c:81
    {
        const double a = 1.0 - (double)0.01f;                    // avg' = a * avg + 0.01f * x (iqconverter_float.c:461-464)
        f->A = 1.0;
        f->narrow = (flags & SDRX_FE_BESIDE_ENGINE) != 0;
        for (int i = 0; i < (f->narrow ? kNSeg : kFeSeg); ++i) f->A *= a;
    }
    std::vector<float> k((size_t)len);
    for (int i = 0, j = 0; i < len; ++i, j += 2) k[i] = hb_kernel[j];   // iqconverter_float.c:91-94
    auto bail = [&](const char *what) { std::string keep = what; sdrx_airspy_fe_destroy(f); return fail(SDRX_ERR_CUDA, keep.c_str()); };
    if (cudaStreamCreateWithFlags(&f->st, cudaStreamNonBlocking) != cudaSuccess) return bail("cudaStreamCreate");
    if (cudaEventCreateWithFlags(&f->ev_done, cudaEventDisableTiming) != cudaSuccess) return bail("cudaEventCreate");
    const size_t pairs = max_samples / 2;
    if (cudaMalloc((void **)&f->d_raw, (size_t)max_samples * 2 + 64) != cudaSuccess ||
        cudaMalloc((void **)&f->d_seg_b, sizeof(float) * (max_samples / kFeSeg + 2)) != cudaSuccess ||
        cudaMalloc((void **)&f->d_seg_start, sizeof(float) * (max_samples / kFeSeg + 2)) != cudaSuccess ||
        cudaMalloc((void **)&f->d_ev, sizeof(float) * (kFeEvenHist + pairs + 8)) != cudaSuccess ||
        cudaMalloc((void **)&f->d_od, sizeof(float) * (kFeOddHist + pairs + 8)) != cudaSuccess ||
        cudaMalloc((void **)&f->d_k, sizeof(float) * len) != cudaSuccess ||
        cudaMalloc((void **)&f->d_avg, sizeof(float) * 2) != cudaSuccess ||
        cudaMalloc((void **)&f->d_iq[0], sizeof(float2) * pairs) != cudaSuccess ||
        cudaMalloc((void **)&f->d_iq[1], sizeof(float2) * pairs) != cudaSuccess)
        return bail("cudaMalloc failed");
    if (cudaMemcpy(f->d_k, k.data(), sizeof(float) * len, cudaMemcpyHostToDevice) != cudaSuccess) return bail("cudaMemcpy failed");
    {   // same (maximum) shared-memory carve-out as the channel bank: an SM only co-schedules kernels that agree on it
        const int mx = (int)cudaSharedmemCarveoutMaxShared;
        cudaFuncSetAttribute(k_fe_dc<true>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_dc<false>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_iq, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_carry, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_partial<true>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_partial<false>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_scan, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        const int nsm = (int)fe_narrow_smem();
        cudaFuncSetAttribute(k_fe_walk_narrow<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, nsm);
        cudaFuncSetAttribute(k_fe_walk_narrow<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, nsm);
        cudaFuncSetAttribute(k_fe_walk_narrow<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, nsm);
        cudaFuncSetAttribute(k_fe_walk_narrow<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, nsm);
        cudaFuncSetAttribute(k_fe_walk_narrow<true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_walk_narrow<true, false>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_walk_narrow<false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaFuncSetAttribute(k_fe_walk_narrow<false, false>, cudaFuncAttributePreferredSharedMemoryCarveout, mx);
        cudaGetLastError();
    }
    int rc = sdrx_airspy_fe_reset(f);
    if (rc) { sdrx_airspy_fe_destroy(f); return rc; }
    *out = f;
    return SDRX_OK;
}

namespace {

int fe_run(sdrx_airspy_fe *f, const void *d_raw, unsigned n) {
    const bool packed = (f->flags & SDRX_FE_PACKED12) != 0;
    const int exact = (f->flags & SDRX_FE_EXACT_DC) ? 1 : 0;
    const int cur = (int)(f->calls & 1);
    const int segs = exact ? 1 : div_up(n, kFeSeg);
    const int npair = (int)(n / 2);
    // Measured on B200 (profiles/r01v_frontend_narrow.txt): confining the passes to 4 CTAs of 512 threads with
    // 512-sample segments -- so that they would fit into the SMs a 143-CTA bank launch leaves free -- makes every
    // recurrence step ~200 cycles (40 warps of 1 KB-strided 16 B loads, nothing to hide their latency behind):
    // 365 us per block instead of 43.  The wide grid stays; the passes keep their grid-stride form.
    auto grid = [](int tasks) { return std::max(1, std::min(kFeMaxCtas, div_up(tasks, kFeThreads))); };
    if (f->narrow && !exact) {
        const int nsegs = div_up(n, kNSeg), whole = (int)(n / kNSeg);
        const size_t sm = fe_narrow_smem();
        if (nsegs > kNBack + 1) {
            if (packed) k_fe_walk_narrow<true, false><<<div_up(whole, kNThreads), kNThreads, sm, f->st>>>(d_raw, (int)n, nullptr, nullptr, 0.0f, nullptr, nullptr, f->d_seg_b);
            else        k_fe_walk_narrow<false, false><<<div_up(whole, kNThreads), kNThreads, sm, f->st>>>(d_raw, (int)n, nullptr, nullptr, 0.0f, nullptr, nullptr, f->d_seg_b);
            k_fe_scan<<<1, 128, 0, f->st>>>(f->d_seg_b, whole, f->d_avg, (float)f->A, f->d_seg_start);
            f->launches += 2;
        }
        if (packed) k_fe_walk_narrow<true, true><<<div_up(nsegs, kNThreads), kNThreads, sm, f->st>>>(d_raw, (int)n, f->d_seg_start, f->d_avg, f->hbc, f->d_ev, f->d_od, nullptr);
        else        k_fe_walk_narrow<false, true><<<div_up(nsegs, kNThreads), kNThreads, sm, f->st>>>(d_raw, (int)n, f->d_seg_start, f->d_avg, f->hbc, f->d_ev, f->d_od, nullptr);
    } else {
    if (!exact && segs > kFeBack + 1) {
            const int whole = (int)(n / kFeSeg);
            if (packed) k_fe_partial<true><<<grid(whole), kFeThreads, 0, f->st>>>(d_raw, (int)n, f->d_seg_b);
            else        k_fe_partial<false><<<grid(whole), kFeThreads, 0, f->st>>>(d_raw, (int)n, f->d_seg_b);
            k_fe_scan<<<1, 128, 0, f->st>>>(f->d_seg_b, whole, f->d_avg, (float)f->A, f->d_seg_start);
            f->launches += 2;
        }
        if (packed) k_fe_dc<true><<<grid(segs), kFeThreads, 0, f->st>>>(d_raw, f->d_seg_start, (int)n, f->d_avg, f->hbc, f->d_ev, f->d_od, exact);
        else        k_fe_dc<false><<<grid(segs), kFeThreads, 0, f->st>>>(d_raw, f->d_seg_start, (int)n, f->d_avg, f->hbc, f->d_ev, f->d_od, exact);
    }
    if (f->out_free) {       // everything up to here only touched the front end's own scratch; the output ring slot is the consumer's
        CU(cudaStreamWaitEvent(f->st, f->out_free, 0));
        f->out_free = nullptr;
    }
    k_fe_iq<<<grid(div_up(npair, 8)), kFeThreads, 0, f->st>>>(f->d_ev, f->d_od, npair, f->d_k, f->len, f->d_iq[cur]);
    k_fe_carry<<<1, 64, 0, f->st>>>((int)n, f->d_ev, f->d_od, f->d_avg);
    CU(cudaGetLastError());
    CU(cudaEventRecord(f->ev_done, f->st));
    f->launches += 3;
    f->total += n;
    ++f->calls;
    f->last_pairs = (unsigned)npair;
    return SDRX_OK;
}

int fe_check(sdrx_airspy_fe *f, const void *raw, unsigned n) {
    if (!f || (!raw && n)) return fail(SDRX_ERR_ARG, "null argument");
    if (n == 0 || n % 8) return fail(SDRX_ERR_ARG, "n_samples must be a positive multiple of 8 (libairspy buffers are)");
    if (n > f->cap) return fail(SDRX_ERR_ARG, "more samples than max_samples");
    return SDRX_OK;
}

}  // namespace

extern "C" int sdrx_airspy_fe_process(sdrx_airspy_fe *f, const void *raw, unsigned n_samples, float *iq_host) {
    int rc = fe_check(f, raw, n_samples);
    if (rc) return rc;
    DevGuard g(f->dev);
    const size_t bytes = (f->flags & SDRX_FE_PACKED12) ? (size_t)n_samples / 8 * 12 : (size_t)n_samples * 2;
    cudaPointerAttributes attr{};
    const bool pinned = cudaPointerGetAttributes(&attr, raw) == cudaSuccess &&
                        (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    const void *from = raw;
    if (!pinned) {
        if (!f->h_stage) CU(cudaMallocHost((void **)&f->h_stage, (size_t)f->cap * 2));
        CU(cudaStreamSynchronize(f->st));           // the previous call's copy out of the staging buffer has run
        memcpy(f->h_stage, raw, bytes);
        from = f->h_stage;
    }
    CU(cudaMemcpyAsync(f->d_raw, from, bytes, cudaMemcpyHostToDevice, f->st));
    rc = fe_run(f, f->d_raw, n_samples);
    if (rc) return rc;
    if (iq_host) {
        CU(cudaMemcpyAsync(iq_host, f->d_iq[(f->calls - 1) & 1], sizeof(float2) * (size_t)(n_samples / 2), cudaMemcpyDeviceToHost, f->st));
        CU(cudaStreamSynchronize(f->st));
    }
    return SDRX_OK;
}

extern "C" int sdrx_airspy_fe_process_device(sdrx_airspy_fe *f, const void *d_raw, unsigned n_samples) {
    int rc = fe_check(f, d_raw, n_samples);
    if (rc) return rc;
    DevGuard g(f->dev);
    return fe_run(f, d_raw, n_samples);
}

extern "C" const float *sdrx_airspy_fe_output(sdrx_airspy_fe *f, unsigned *n_pairs) {
    if (!f) return nullptr;
    if (n_pairs) *n_pairs = f->last_pairs;
    return reinterpret_cast<const float *>(f->d_iq[(f->calls + 1) & 1]);    // the buffer the last call wrote
}
// Orders the NEXT call's output write behind an event of the consumer (e.g. sdrx_engine_ingest_event after the push
// that read the previous output): the output ring is two deep, so waiting for the consumer of call k-1 frees call
// k+1's slot.  Only the last pass writes the ring, so the copy in and the remove_dc passes of call k+1 still run ahead.
extern "C" int sdrx_airspy_fe_wait_event(sdrx_airspy_fe *f, void *cuda_event) {
    if (!f || !cuda_event) return fail(SDRX_ERR_ARG, "null argument");
    f->out_free = (cudaEvent_t)cuda_event;
    return SDRX_OK;
}
extern "C" void *sdrx_airspy_fe_done_event(sdrx_airspy_fe *f) { return f ? (void *)f->ev_done : nullptr; }
extern "C" void *sdrx_airspy_fe_stream(sdrx_airspy_fe *f) { return f ? (void *)f->st : nullptr; }
extern "C" uint64_t sdrx_airspy_fe_launch_count(sdrx_airspy_fe *f) { return f ? f->launches : 0; }
extern "C" int sdrx_airspy_fe_sync(sdrx_airspy_fe *f) {
    if (!f) return fail(SDRX_ERR_ARG, "null front end");
    DevGuard g(f->dev);
    CU(cudaStreamSynchronize(f->st));
    return SDRX_OK;
}
extern "C" int sdrx_airspy_fe_avg(sdrx_airspy_fe *f, float *avg) {
    if (!f || !avg) return fail(SDRX_ERR_ARG, "null argument");
    DevGuard g(f->dev);
    CU(cudaStreamSynchronize(f->st));
    CU(cudaMemcpy(avg, f->d_avg, sizeof(float), cudaMemcpyDeviceToHost));
    return SDRX_OK;
}

// Debug: see g_fe_trace.  enable allocates / clears 10 u64 and points the kernels at them; read copies them out.
static unsigned long long *g_fe_trace_buf = nullptr;
extern "C" int sdrx_debug_fe_trace(int enable) {
    unsigned long long *p = nullptr;
    if (enable) {
        if (!g_fe_trace_buf) CU(cudaMalloc((void **)&g_fe_trace_buf, sizeof(unsigned long long) * 10));
        CU(cudaMemset(g_fe_trace_buf, 0, sizeof(unsigned long long) * 10));
        p = g_fe_trace_buf;
    }
    CU(cudaMemcpyToSymbol(g_fe_trace, &p, sizeof(p)));
    return SDRX_OK;
}
extern "C" int sdrx_debug_fe_trace_read(unsigned long long *out) {
    if (!g_fe_trace_buf || !out) return 0;
    return cudaMemcpy(out, g_fe_trace_buf, sizeof(unsigned long long) * 10, cudaMemcpyDeviceToHost) == cudaSuccess ? 10 : 0;
}

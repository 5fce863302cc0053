// sdrx_b200 -- the table GEMM of tgemm.cuh on the 5th-generation tensor cores (tcgen05.mma, accumulators in tensor memory,
// operands staged by TMA): north_star's "tensor cores only if the mix-plus-FIR is recast as a complex GEMM (channels x taps),
// and only if that version stays within tolerance".  SDRX_FLAG_TGEMM_TC.
//
// The product is the same as k_tgemm's -- per pair slot s and output instant M, over the table period p < N:
//     P = sum Tr[s][p] Ur[p][M],  R = sum Tr Ui,  S = sum Ti Ur,  Q = sum Ti Ui;   Y_first = (P - Q, R + S),  Y_partner = (P + Q, R - S)
// (the reference's own translator tables, sdrx.cpp:1682-1690, are the A matrix; msd.hpp:133-154, :251-335 is the map) -- written as
// ONE real GEMM  D[2 slots x 2 instants] = A[rows = (Tr_s, Ti_s)][p] * B[cols = (Ur_M, Ui_M)][p]^T.
//
// Precision.  A tensor-core product takes bf16 (8 significant bits) operands; an FP32 value is EXACTLY the sum of three bf16
// values (x = x0 + x1 + x2, each the round-to-nearest bf16 of what is left), so a x b = sum over the nine plane products
// a_i x b_j, every one of them exact in the FP32 accumulator's product (8 x 8 bits).  kProducts = 9 keeps them all (the
// products are then EXACT and the only rounding left is the accumulation); kProducts = 6 (the default) drops the three
// smallest (a1 b2, a2 b1, a2 b2: < 2^-23 of the term together; measured accuracy identical).
// Accumulation is the accuracy lever.  A single running sum over the 1200 terms loses weak channels, and so does rounding P and
// Q -- each as large as the STRONGER channel of a +-offset pair -- before the weaker one is formed as their difference.  So
// every 16 table entries (one k-step) get their own tensor-memory accumulator (longer units put more of the tensor core's own
// accumulate rounding into the result), the epilogue warps add the finished units into FP32 register totals with Kahan
// compensation while the tensor core works on the following ones, and the pair is un-mixed from both the totals and the
// compensation terms.  Measured on config 3: worst channel IQ 2.9e-5 against the oracle (1.36e-4 without compensation).
//
//   k_fold_tc<FMT>  the fold of tgemm.cuh (by default in double, rounded once to FP32), written as the three bf16 planes of B,
//                   k-major: B[stream][plane][2 M + (re|im)][p]; four 128-thread CTAs per output instant, on the engine's prep
//                   stream under the previous block's GEMM (B planes ping-pong)
//   k_tgemm_tc      one CTA per 128-row x 64-column tile of D (64 pair slots x 32 instants), 10 warps:
//                     warp 8  TMA producer: per 64-entry k-block the three planes of A (128 x 64 bf16) and of B (64 x 64 bf16)
//                             through 2-D tensor maps with the 128-byte swizzle, a three-deep ring of 72 KB stages on mbarriers
//                     warp 9  allocates tensor memory; from a warp-uniform loop one elected lane issues tcgen05.mma (M 128, N 64,
//                             K 16, kind::f16 with bf16 inputs, FP32 accumulate) -- four k-steps x kProducts per k-block into a
//                             GROUP of four accumulators -- and ONE tcgen05.commit per k-block (a commit drains the tensor pipe:
//                             ~300 cycles each); two groups = eight 64-column accumulators = all 512 columns
//                     warps 0-7  epilogue: tcgen05.ld of the group's accumulators (32 lanes x 32 columns per warp: a warp reaches
//                             only its own lane quadrant, two warps share a quadrant's columns), Kahan add into the totals; at
//                             the end un-mix the pair (neighbouring lanes hold the Tr and Ti rows of a slot) and store
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include "tgemm.cuh"

namespace sdrx {

constexpr int kTcRows = 128;                  // UMMA M
constexpr int kTcCols = 64;                   // UMMA N
constexpr int kTcK = 64;                      // bf16 elements of one staged k-block row = 128 bytes = one swizzle row
constexpr int kTcKStep = 16;                  // K of one tcgen05.mma with 16-bit inputs
constexpr int kTcStages = 3;                  // at most; the launch picks 2 or 3 (TcGemm::stages)
constexpr int kTcPlanes = 3;
constexpr int kTcAcc = 8;                     // accumulators in the tensor-memory ring
constexpr int kTcTmemCols = kTcAcc * kTcCols; // 512: all of it
constexpr int kTcEpiWarps = 8;                // epilogue warps: warp w drains TMEM lanes 32 (w % 4) .. + 31, columns 32 (w / 4) .. + 31 of an accumulator
constexpr int kTcThreads = (kTcEpiWarps + 2) * 32;
constexpr int kTcABytes = kTcRows * kTcK * 2; // 16 KB per plane
constexpr int kTcBBytes = kTcCols * kTcK * 2; //  8 KB per plane
constexpr int kTcStageBytes = kTcPlanes * (kTcABytes + kTcBBytes);              // 72 KB
inline size_t tc_smem_bytes(int stages) { return (size_t)stages * kTcStageBytes + 1024 + 256; }   // + alignment slack + barriers / tmem slot
constexpr size_t kTcSmem = (size_t)kTcStages * kTcStageBytes + 1024 + 256;

// ---- tcgen05 / tensor-memory primitives (inline PTX; SASS: UTCHMMA, LDTM, UTCBAR, UTMALDG)
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_alloc(uint32_t *slot, unsigned cols) {        // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_dealloc(uint32_t taddr, unsigned cols) {      // whole warp, the one that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T; one thread issues
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on `bar` once every tcgen05.mma issued before it by this thread has completed (implies fence::before_thread_sync)
__device__ __forceinline__ void tc_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// ... the same arrival on the barrier at this offset in every CTA of the cluster named by `mask`
__device__ __forceinline__ void tc_commit_mc(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// one box of a 2-D tensor map into the shared memory of every CTA of the cluster named by `mask` (same offset in each),
// completing its bytes on the barrier at `bar`'s offset in each of them
__device__ __forceinline__ void tma_load_2d_mc(void *dst, const void *tmap, int x, int y, uint64_t *bar, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%2, %3}], [%4], %5;"
                 ::"r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ unsigned cluster_rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// one lane of the (converged) warp, chosen by the hardware: the warp-uniform way to issue single-thread instructions
__device__ __forceinline__ bool tc_elect_one() {
    uint32_t p;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(p));
    return p != 0;
}
// 32 lanes x 32 consecutive columns: thread t of the warp gets lane (warp % 4) * 32 + t
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t *v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, "
        "%21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]),
          "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]),
          "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor of a K-major operand tile whose rows are 128 bytes (64 bf16) and whose 8-row groups are
// 1024-byte atoms under the 128-byte swizzle -- exactly what a {64, rows} TMA box with CU_TENSOR_MAP_SWIZZLE_128B writes.
// bits [0,14) address >> 4, [16,30) leading byte offset >> 4 (unused for this layout: 1, as CUTLASS sets it), [32,46) stride byte offset >> 4 (one
// atom = 1024 B), [46,48) version 1 (sm_100), [61,64) layout type 2 (SWIZZLE_128B).  A k-step of 16 elements = 32 bytes
// further along the row is the same descriptor with the address advanced by 32 bytes (+2 in the field).
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// Instruction descriptor, kind::f16: [4,6) D format 1 = F32, [7,10) A format 1 = BF16, [10,13) B format 1 = BF16, bit 15 / 16 A / B
// major 0 = K, [17,23) N >> 3, [24,29) M >> 4.
constexpr uint32_t kTcIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTcCols >> 3) << 17) | ((uint32_t)(kTcRows >> 4) << 24);

__device__ __forceinline__ void bf16_split3(float x, __nv_bfloat16 &h0, __nv_bfloat16 &h1, __nv_bfloat16 &h2) {
    h0 = __float2bfloat16_rn(x);
    const float r1 = __fsub_rn(x, __bfloat162float(h0));   // exact
    h1 = __float2bfloat16_rn(r1);
    const float r2 = __fsub_rn(r1, __bfloat162float(h1));  // exact
    h2 = __float2bfloat16_rn(r2);                          // exact: at most 8 significant bits are left
}

// The fold of tgemm.cuh (k_fold) writing B: [stream][plane][2 M + (re | im)][p] bf16, p contiguous (pitch Kp elements).
// grid (ceil(n_out / kFoldTcM), n_streams).  Small CTAs, one output instant each: the fold runs under the previous block's GEMM, whose
// CTAs leave room for ~420 more threads per SM -- 128-thread CTAs fit three beside each (one wave), 256-thread CTAs fit one (two
// waves: 55 us in the pipeline for 17 us of work, and the next GEMM waits for the fold).
constexpr int kFoldTcM = 1, kFoldTcThreads = 128;
constexpr int kFoldTcSplit = 4;               // CTAs per output instant (blockIdx.z: a quarter of the table period each): 2.3 items per thread
template <int FMT>
static __global__ void __launch_bounds__(kFoldTcThreads)
k_fold_tc(const void *__restrict__ raw_all, long stream_bytes, int hist, int valid_lo, int pos0, int D, int N, const double *__restrict__ hc, int Lh,
          int n_out, __nv_bfloat16 *__restrict__ B, int cols, int Kp, const float *__restrict__ hc32, int dbl) {
    const void *raw = reinterpret_cast<const unsigned char *>(raw_all) + (long)blockIdx.y * stream_bytes;
    __nv_bfloat16 *Bs = B + (size_t)blockIdx.y * kTcPlanes * cols * Kp;
    const size_t plane = (size_t)cols * Kp;
    const int M0 = blockIdx.x * kFoldTcM;
    const int per = (kFoldTcM * N + kFoldTcSplit - 1) / kFoldTcSplit;
    const int it_end = min(kFoldTcM * N, (int)(blockIdx.z + 1) * per);
    for (int it = blockIdx.z * per + threadIdx.x; it < it_end; it += kFoldTcThreads) {
        const int m = it / N, p = it - m * N;
        const int M = M0 + m;
        if (M >= n_out) break;
        const int nM = D * M + D - 1;
        int d = (pos0 + nM - p) % N;
        if (d < 0) d += N;
        // the 7-8 terms of a bin.  kFoldDouble: double taps (the exact convolution of the reference's float stage taps) and a double
        // sum, rounded once to FP32; otherwise float taps and an FP32 fma chain (half the tap bytes through L2, no FP64: step -3 %).
        // Measured against the oracle: worst channel IQ 2.89e-5 / 2.97e-5, median 0.97e-5 / 1.18e-5 -- the accuracy lever is the
        // compensated accumulation of the GEMM, the fold in double only polishes the median
        float2 acc;
        if (dbl) {
            double ar = 0.0, ai = 0.0;
            for (int j0 = d; j0 < Lh; j0 += 8 * N) {
                float2 xv[8];
                double hv[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int j = j0 + q * N;
                    const bool in = j < Lh;
                    xv[q] = in ? load_iq_from<FMT>(raw, (long)hist + nM - j, valid_lo) : make_float2(0.f, 0.f);
                    hv[q] = in ? __ldg(hc + j) : 0.0;
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    ar = fma(hv[q], (double)xv[q].x, ar);
                    ai = fma(hv[q], (double)xv[q].y, ai);
                }
            }
            acc = make_float2(__double2float_rn(ar), __double2float_rn(ai));
        } else {
            acc = make_float2(0.f, 0.f);
            for (int j0 = d; j0 < Lh; j0 += 8 * N) {
                float2 xv[8];
                float hv[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const int j = j0 + q * N;
                    const bool in = j < Lh;
                    xv[q] = in ? load_iq_from<FMT>(raw, (long)hist + nM - j, valid_lo) : make_float2(0.f, 0.f);
                    hv[q] = in ? __ldg(hc32 + j) : 0.0f;
                }
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    acc.x = fmaf(hv[q], xv[q].x, acc.x);
                    acc.y = fmaf(hv[q], xv[q].y, acc.y);
                }
            }
        }
        __nv_bfloat16 r0, r1, r2, i0, i1, i2;
        bf16_split3(acc.x, r0, r1, r2);
        bf16_split3(acc.y, i0, i1, i2);
        __nv_bfloat16 *re = Bs + (size_t)(2 * M) * Kp + p, *im = re + Kp;
        re[0] = r0; re[plane] = r1; re[2 * plane] = r2;
        im[0] = i0; im[plane] = i1; im[2 * plane] = i2;
    }
}

// grid (rows_pad / 128, cols / 64, n_streams), clusters of (1, cl, 1) CTAs (cl = 1: none); A map box {64, 128 / cl}; A map: [3 rows_pad][N] bf16, box {64, 128}; B map: [n_streams 3 cols][N] bf16, box {64, 64}.
// kProducts: 9 (all plane products: exact) or 6; kSteps: 16-entry k-steps per drain unit (1, 2 or 4).
template <int kProducts, int kSteps>
static __global__ void __launch_bounds__(kTcThreads, 1)
k_tgemm_tc(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, int rows_pad, int cols, int N,
           const int *__restrict__ slot_chan, int chans_per_stream, float2 *__restrict__ ch_iq, long iq_stride, int n_out, int cl, int dbg, unsigned long long *__restrict__ trace) {
    const bool comp = !(dbg & 4);                                 // SDRX_TC_DEBUG=4: plain FP32 adds of the unit sums (experiment)
    extern __shared__ unsigned char tc_smem_raw[];
    unsigned char *sm = reinterpret_cast<unsigned char *>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);   // swizzle atoms: 1024-byte aligned
    const int n_st = (dbg & 8) ? 2 : kTcStages;                  // operand stages in the ring (bit 3 of the knob word: two)
    uint64_t *bars = reinterpret_cast<uint64_t *>(sm + (size_t)n_st * kTcStageBytes);
    uint64_t *full = bars, *empty = bars + kTcStages, *tfull = bars + 2 * kTcStages, *tempty = tfull + kTcAcc;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + kTcAcc);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_kb = (N + kTcK - 1) / kTcK;                       // staged k-blocks (the planes are zero past N)
    // SDRX_TC_DEBUG bit 4: per-role cycle counts (total, spent waiting) of every CTA into `trace`
    long long t_wait = 0, t_wait2 = 0, t_sec = 0, t_sec2 = 0;
    const long long t_begin = trace ? clock64() : 0;
#define TC_TIMED_WAIT(acc_, bar_, par_) do { if (trace) { const long long w0_ = clock64(); mbar_wait(bar_, par_); acc_ += clock64() - w0_; } else mbar_wait(bar_, par_); } while (0)

    if (threadIdx.x == 0) {
        for (int i = 0; i < kTcStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, (unsigned)cl); }
        for (int i = 0; i < kTcAcc; ++i) { mbar_init(tfull + i, 1); mbar_init(tempty + i, kTcEpiWarps); }   // [0..1] used: one pair per accumulator group
        mbar_fence_init();
    }
    if (warp == kTcEpiWarps + 1) tc_alloc(tmem_slot, kTcTmemCols);
    tc_fence_before();
    __syncthreads();
    if (cl > 1) cluster_sync_all();                               // every CTA's barriers exist before a neighbour's copy or commit lands on them
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint16_t mc_mask = (uint16_t)((1u << cl) - 1u);

    if (warp == kTcEpiWarps) {
        // ===== TMA producer: the whole warp walks the loop, one elected lane issues (no divergent region around the bulk copies)
        const int row0 = blockIdx.x * kTcRows;
        const int col0 = blockIdx.z * kTcPlanes * cols + blockIdx.y * kTcCols;
        int st = 0;
        unsigned ph = 1;                                          // parity of "the consumer has freed the stage": a fresh barrier passes
        for (int kb = 0; kb < n_kb; ++kb) {
            TC_TIMED_WAIT(t_wait, empty + st, ph);
            if (tc_elect_one()) {
                unsigned char *a = sm + (size_t)st * kTcStageBytes, *b = a + kTcPlanes * kTcABytes;
                if (dbg & 2) mbar_arrive(full + st);              // timing experiment (SDRX_TC_DEBUG=2): no operand loads, wrong results
                else {
                    mbar_expect_tx(full + st, (unsigned)kTcStageBytes);
                    if (cl > 1) {
                        // the CTAs of a cluster share the row tile: each fetches 1 / cl of every A plane and multicasts it to all of them
                        const int rank = (int)cluster_rank(), slice = kTcRows / cl;
#pragma unroll
                        for (int pl = 0; pl < kTcPlanes; ++pl) {
                            tma_load_2d_mc(a + pl * kTcABytes + rank * slice * (kTcK * 2), &map_a, kb * kTcK, pl * rows_pad + row0 + rank * slice, full + st, mc_mask);
                            tma_load_2d(b + pl * kTcBBytes, &map_b, kb * kTcK, pl * cols + col0, full + st);
                        }
                    } else {
#pragma unroll
                        for (int pl = 0; pl < kTcPlanes; ++pl) {
                            tma_load_2d(a + pl * kTcABytes, &map_a, kb * kTcK, pl * rows_pad + row0, full + st);
                            tma_load_2d(b + pl * kTcBBytes, &map_b, kb * kTcK, pl * cols + col0, full + st);
                        }
                    }
                }
            }
            __syncwarp();
            if (++st == n_st) { st = 0; ph ^= 1u; }
        }
    } else if (warp == kTcEpiWarps + 1) {
        // ===== MMA issuer: warp-uniform loop, one elected lane issues.  Per staged k-block four k-steps of 16 table entries; a drain
        // unit is kSteps of them (kSteps divides 4, so units never straddle a stage).
        constexpr int pa[9] = {2, 2, 1, 2, 0, 1, 1, 0, 0};      // plane products, smallest first: (a plane, b plane)
        constexpr int pb[9] = {2, 1, 2, 0, 2, 1, 0, 1, 0};
        constexpr int first_prod = 9 - kProducts;                 // 6 products: skip (2,2), (2,1), (1,2)
        // Accumulators are handed over in GROUPS of four (one staged k-block = four k-steps): a tcgen05.commit costs the tensor pipe
        // ~300 cycles of drain (measured: the fixed cost per commit), so there is one per k-block, not one per 16-entry unit -- the
        // units stay 16 entries long (that is what bounds the rounding inside the tensor core), each in its own accumulator.
        int st = 0, grp = 0;
        unsigned ph = 0, gph = 1;                                 // stage-full parity; group-free parity (fresh barriers pass)
        for (int kb = 0; kb < n_kb; ++kb) {
            TC_TIMED_WAIT(t_wait, full + st, ph);
            TC_TIMED_WAIT(t_wait2, tempty + grp, gph);
            tc_fence_after();
            const uint64_t da = tc_smem_desc(smem_u32(sm + (size_t)st * kTcStageBytes));
            const uint64_t db = tc_smem_desc(smem_u32(sm + (size_t)st * kTcStageBytes + kTcPlanes * kTcABytes));
            const long long s0_ = trace ? clock64() : 0;
            if (tc_elect_one()) {
#pragma unroll
                for (int sub = 0; sub < 4; ++sub) {
                    const bool unit_first = (sub % kSteps) == 0;
                    const uint32_t d_tmem = tmem + (uint32_t)((grp * 4 + sub / kSteps) * kTcCols);
#pragma unroll
                    for (int q = first_prod; q < 9; ++q)
                        if (!(dbg & 1) || q == 8)                 // timing experiment (SDRX_TC_DEBUG=1): one product instead of nine, wrong results
                            tc_mma_bf16(d_tmem, da + (uint64_t)((pa[q] * kTcABytes + sub * (kTcKStep * 2)) >> 4),
                                        db + (uint64_t)((pb[q] * kTcBBytes + sub * (kTcKStep * 2)) >> 4), kTcIdesc,
                                        (unit_first && (q == first_prod || (dbg & 1))) ? 0u : 1u);
                }
                if (cl > 1) tc_commit_mc(empty + st, mc_mask); else tc_commit(empty + st);   // the stage is free once these have read it
                tc_commit(tfull + grp);                                                       // ... and the group's accumulators are final
            }
            __syncwarp();
            if (trace) t_sec += clock64() - s0_;
            if (++grp == 2) { grp = 0; gph ^= 1u; }
            if (++st == n_st) { st = 0; ph ^= 1u; }
        }
    } else {
        // ===== epilogue warps 0..7: warp w owns TMEM lanes 32 (w % 4) .. + 31 (the rows of the tile; a warp can only reach its own
        // lane quadrant) and columns 32 (w / 4) .. + 31 of every accumulator.
        // Unit sums are added with Kahan compensation: tot - cmp is the sum of the units to ~48 bits, so what is left of FP32
        // rounding is what the tensor core does inside a unit (16 entries).  It matters because of the pairing: P and Q are each
        // as large as the STRONGER channel of a +-offset pair, and the weaker one is their small difference.
        constexpr int kEc = kTcCols * 4 / kTcEpiWarps;            // columns per warp: 32
        static_assert(kEc == 32, "the drain is one 32-column tcgen05.ld per warp");
        const int quad = warp & 3, chalf = warp >> 2;
        float tot[kEc], cmp[kEc];
#pragma unroll
        for (int i = 0; i < kEc; ++i) tot[i] = cmp[i] = 0.0f;
        const uint32_t tlane = tmem + ((uint32_t)(quad * 32) << 16) + chalf * kEc;
        int grp = 0;
        unsigned gph = 0;
        for (int kb = 0; kb < n_kb; ++kb) {
            TC_TIMED_WAIT(t_wait, tfull + grp, gph);
            tc_fence_after();
            constexpr int kUnits = 4 / kSteps;                    // accumulators of a group in use
            constexpr int kBatch = kUnits >= 2 ? 2 : 1;           // accumulators read per tcgen05.wait::ld (register budget)
#pragma unroll
            for (int u0 = 0; u0 < kUnits; u0 += kBatch) {
                uint32_t v[kBatch][kEc];
#pragma unroll
                for (int u = 0; u < kBatch; ++u) tc_ld32(tlane + (uint32_t)((grp * 4 + u0 + u) * kTcCols), v[u]);
                tc_ld_wait();
                if (u0 + kBatch == kUnits) {
                    tc_fence_before();
                    if (lane == 0) mbar_arrive(tempty + grp);     // this warp's part of the group is in registers
                }
#pragma unroll
                for (int u = 0; u < kBatch; ++u)                  // unit by unit, in k order
#pragma unroll
                    for (int j = 0; j < kEc; ++j) {
                        if (comp) {
                            const float y = __fsub_rn(__uint_as_float(v[u][j]), cmp[j]);
                            const float t = __fadd_rn(tot[j], y);
                            cmp[j] = __fsub_rn(__fsub_rn(t, tot[j]), y);
                            tot[j] = t;
                        } else {
                            tot[j] = __fadd_rn(tot[j], __uint_as_float(v[u][j]));
                        }
                    }
            }
            if (++grp == 2) { grp = 0; gph ^= 1u; }
        }
        // un-mix: even row = Tr of the slot -> (P, R) per instant, odd row = Ti -> (S, Q); the partner row is the neighbouring lane
        const int row = blockIdx.x * kTcRows + quad * 32 + lane;
        const int odd = row & 1;
        const int c = slot_chan[row];                              // [slot][2]: even row -> first channel, odd row -> conjugate partner
        const int m0 = blockIdx.y * (kTcCols / 2) + chalf * (kEc / 2);
        float2 *dst = c >= 0 ? ch_iq + ((long)blockIdx.z * chans_per_stream + c) * iq_stride + m0 : nullptr;
#pragma unroll
        for (int m = 0; m < kEc / 2; m += 2) {
            float4 y;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float a = tot[2 * (m + h)], b = tot[2 * (m + h) + 1];
                const float oa = __shfl_xor_sync(0xffffffffu, a, 1), ob = __shfl_xor_sync(0xffffffffu, b, 1);
                // even: a = P, b = R, oa = S, ob = Q -> (P - Q, R + S);  odd: a = S, b = Q, oa = P, ob = R -> (P + Q, R - S)
                float re = odd ? __fadd_rn(oa, b) : __fsub_rn(a, ob);
                float im = odd ? __fsub_rn(ob, a) : __fadd_rn(b, oa);
                if (comp) {                                        // the low parts: every total is tot - cmp
                    const float ca = cmp[2 * (m + h)], cb = cmp[2 * (m + h) + 1];
                    const float oca = __shfl_xor_sync(0xffffffffu, ca, 1), ocb = __shfl_xor_sync(0xffffffffu, cb, 1);
                    re = __fsub_rn(re, odd ? __fadd_rn(oca, cb) : __fsub_rn(ca, ocb));
                    im = __fsub_rn(im, odd ? __fsub_rn(ocb, ca) : __fadd_rn(cb, oca));
                }
                if (h == 0) { y.x = re; y.y = im; } else { y.z = re; y.w = im; }
            }
            if (dst && m0 + m < n_out) *reinterpret_cast<float4 *>(dst + m) = y;   // n_out is even (tgemm_eligible)
        }
    }
    if (trace && lane == 0 && (warp == 0 || warp >= kTcEpiWarps)) {
        const int cta = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x, role = warp == 0 ? 0 : warp - kTcEpiWarps + 1;   // 0 epilogue, 1 producer, 2 issuer
        unsigned long long *o = trace + ((size_t)cta * 3 + role) * 8;
        o[0] = (unsigned long long)(clock64() - t_begin); o[1] = (unsigned long long)t_wait; o[2] = (unsigned long long)t_wait2; o[3] = (unsigned long long)t_sec; o[4] = (unsigned long long)t_sec2;
    }
#undef TC_TIMED_WAIT
    tc_fence_before();
    __syncthreads();
    if (cl > 1) cluster_sync_all();                               // no CTA leaves while a neighbour's commit may still arrive on its barriers
    if (warp == kTcEpiWarps + 1) tc_dealloc(tmem, kTcTmemCols);
}

// ---- host side
struct TcGemm {
    bool enabled = false;
    int rows_pad = 0, cols = 0, Kp = 0, products = 9, steps = 1, cluster = 1, debug = 0;
    __nv_bfloat16 *d_a = nullptr;      // [3][rows_pad][Kp]
    __nv_bfloat16 *d_b[2] = {nullptr, nullptr};   // [n_streams][3][cols][Kp], ping-pong: the fold of block b + 1 (copy stream) writes one while
                                                  // the GEMM of block b (main stream) reads the other
    cudaEvent_t ev_b_free[2] = {nullptr, nullptr};  // the GEMM that read d_b[par] has run
    bool b_busy[2] = {false, false};
    double *d_hc64 = nullptr;          // the composite prototype in double (the fold's taps)
    unsigned long long *d_trace = nullptr;   // SDRX_TC_DEBUG bit 4: [ctas][3 roles][4]
    int trace_left = 0;
    int fold_double = 1;               // the fold with double taps and a double sum (SDRX_TC_FOLD_DOUBLE=0: float taps, FP32 fma chain)
    int *d_row_chan = nullptr;         // [rows_pad]
    CUtensorMap map_a, map_b[2];
    void destroy() {
        cudaFree(d_a); cudaFree(d_b[0]); cudaFree(d_b[1]); cudaFree(d_row_chan);
        for (int i = 0; i < 2; ++i) { if (ev_b_free[i]) cudaEventDestroy(ev_b_free[i]); ev_b_free[i] = nullptr; b_busy[i] = false; }
        cudaFree(d_hc64); cudaFree(d_trace); d_trace = nullptr;
        d_a = d_b[0] = d_b[1] = nullptr; d_row_chan = nullptr; d_hc64 = nullptr; enabled = false;
    }
};

inline void host_bf16_split3(float x, uint16_t out[3]) {
    auto rn = [](float v) -> uint16_t {                    // float -> bf16, round to nearest even (finite inputs)
        uint32_t u; memcpy(&u, &v, 4);
        u += 0x7FFFu + ((u >> 16) & 1u);
        return (uint16_t)(u >> 16);
    };
    auto up = [](uint16_t h) -> float { uint32_t u = (uint32_t)h << 16; float v; memcpy(&v, &u, 4); return v; };
    float r = x;
    for (int i = 0; i < 3; ++i) { out[i] = rn(r); r = r - up(out[i]); }
}

template <typename MakeMap>
int tc_init(TcGemm &tc, const TableGemm &tg, const float *tables, const uint8_t *has, const std::vector<int> &slot_chan, int products, int steps,
            int cluster, MakeMap make_map) {
    const int N = tg.N;
    tc.products = products == 6 ? 6 : 9;
    tc.steps = steps == 4 ? 4 : (steps == 2 ? 2 : 1);
    tc.cluster = (cluster == 2 || cluster == 4 || cluster == 8) ? cluster : 1;
    tc.Kp = (N + kTcK - 1) / kTcK * kTcK;                      // row pitch a multiple of 128 bytes: every box row is one aligned L2 line
    tc.rows_pad = (2 * tg.slots + kTcRows - 1) / kTcRows * kTcRows;
    tc.cols = (2 * tg.Mpad + kTcCols - 1) / kTcCols * kTcCols;
    std::vector<uint16_t> a((size_t)kTcPlanes * tc.rows_pad * tc.Kp, 0);
    std::vector<int> row_chan((size_t)tc.rows_pad, -1);
    const size_t plane = (size_t)tc.rows_pad * tc.Kp;
    for (int s = 0; s < tg.slots; ++s) {
        const int c = slot_chan[2 * s];
        row_chan[2 * s] = c;
        row_chan[2 * s + 1] = slot_chan[2 * s + 1];
        for (int p = 0; p < N; ++p) {
            float t[2] = {1.0f, 0.0f};
            if (tables && has[c]) { t[0] = tables[((size_t)c * N + p) * 2]; t[1] = tables[((size_t)c * N + p) * 2 + 1]; }
            for (int ri = 0; ri < 2; ++ri) {
                uint16_t h[3];
                host_bf16_split3(t[ri], h);
                for (int pl = 0; pl < kTcPlanes; ++pl) a[pl * plane + (size_t)(2 * s + ri) * tc.Kp + p] = h[pl];
            }
        }
    }
    CU(cudaMalloc((void **)&tc.d_a, a.size() * 2));
    CU(cudaMemcpy(tc.d_a, a.data(), a.size() * 2, cudaMemcpyHostToDevice));
    const size_t b_elems = (size_t)tg.n_streams * kTcPlanes * tc.cols * tc.Kp;
    for (int i = 0; i < 2; ++i) {
        CU(cudaMalloc((void **)&tc.d_b[i], b_elems * 2));
        CU(cudaMemset(tc.d_b[i], 0, b_elems * 2));
        CU(cudaEventCreateWithFlags(&tc.ev_b_free[i], cudaEventDisableTiming));
    }
    CU(cudaMalloc((void **)&tc.d_hc64, sizeof(double) * tg.h_hc64.size()));
    CU(cudaMemcpy(tc.d_hc64, tg.h_hc64.data(), sizeof(double) * tg.h_hc64.size(), cudaMemcpyHostToDevice));
    CU(cudaMalloc((void **)&tc.d_row_chan, sizeof(int) * row_chan.size()));
    CU(cudaMemcpy(tc.d_row_chan, row_chan.data(), sizeof(int) * row_chan.size(), cudaMemcpyHostToDevice));
    int rc;
    if ((rc = make_map(&tc.map_a, tc.d_a, (uint64_t)tc.Kp, (uint64_t)kTcPlanes * tc.rows_pad, (uint64_t)tc.Kp * 2, kTcK, kTcRows / tc.cluster))) return rc;
    for (int i = 0; i < 2; ++i)
        if ((rc = make_map(&tc.map_b[i], tc.d_b[i], (uint64_t)tc.Kp, (uint64_t)tg.n_streams * kTcPlanes * tc.cols, (uint64_t)tc.Kp * 2, kTcK, kTcCols))) return rc;
    tc.enabled = true;
    return SDRX_OK;
}

inline const void *tc_kernel_ptr(int products, int steps) {
    if (products == 6) return steps == 4 ? (const void *)k_tgemm_tc<6, 4> : steps == 2 ? (const void *)k_tgemm_tc<6, 2> : (const void *)k_tgemm_tc<6, 1>;
    return steps == 4 ? (const void *)k_tgemm_tc<9, 4> : steps == 2 ? (const void *)k_tgemm_tc<9, 2> : (const void *)k_tgemm_tc<9, 1>;
}

// The fold of one block into the B planes d_b[par], on `st` (the engine's copy stream: it runs under the previous block's GEMM).
template <int FMT>
int tgemm_tc_fold(TableGemm &tg, TcGemm &tc, cudaStream_t st, int par, const void *raw, long stream_bytes, int hist, int valid_lo, int pos0, int n_out,
                  uint64_t *launches) {
    if (tc.b_busy[par]) CU(cudaStreamWaitEvent(st, tc.ev_b_free[par], 0));     // the GEMM of two blocks ago read these planes
    k_fold_tc<FMT><<<dim3((n_out + kFoldTcM - 1) / kFoldTcM, tg.n_streams, kFoldTcSplit), kFoldTcThreads, 0, st>>>(raw, stream_bytes, hist, valid_lo, pos0, tg.D, tg.N,
                                                                                              tc.d_hc64, tg.Lh, n_out, tc.d_b[par], tc.cols, tc.Kp, tg.d_hc, tc.fold_double);
    *launches += 1;
    CU(cudaGetLastError());
    return SDRX_OK;
}

// The GEMM over the planes d_b[par], on `st` (the main stream; the caller has made it wait for the fold).
inline int tgemm_tc_launch(TableGemm &tg, TcGemm &tc, cudaStream_t st, int par, int n_out, float2 *ch_iq, long iq_stride, uint64_t *launches,
                           cudaEvent_t ev0 = nullptr, cudaEvent_t ev1 = nullptr) {
    if (ev0) CU(cudaEventRecord(ev0, st));
    const dim3 grid(tc.rows_pad / kTcRows, (2 * n_out + kTcCols - 1) / kTcCols, tg.n_streams);
    const int cl = (tc.cluster > 1 && grid.y % tc.cluster == 0) ? tc.cluster : 1;
    if (cl != tc.cluster) return fail(SDRX_ERR_UNSUPPORTED, "tensor-core table GEMM: the column tiles of this block length do not fill whole clusters");
    cudaLaunchConfig_t lc{};
    lc.gridDim = grid; lc.blockDim = dim3(kTcThreads); lc.dynamicSmemBytes = tc_smem_bytes((tc.debug & 8) ? 2 : kTcStages); lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = (unsigned)cl; at[0].val.clusterDim.z = 1;
    lc.attrs = at; lc.numAttrs = cl > 1 ? 1 : 0;
#define SDRX_TC_GO(P, S)                                                                                                                    \
    CU(cudaLaunchKernelEx(&lc, k_tgemm_tc<P, S>, tc.map_a, tc.map_b[par], tc.rows_pad, tc.cols, tg.N, (const int *)tc.d_row_chan, tg.chans, ch_iq, \
                          iq_stride, n_out, cl, tc.debug, tc.d_trace))
    if (tc.products == 6) { if (tc.steps == 4) SDRX_TC_GO(6, 4); else if (tc.steps == 2) SDRX_TC_GO(6, 2); else SDRX_TC_GO(6, 1); }
    else                  { if (tc.steps == 4) SDRX_TC_GO(9, 4); else if (tc.steps == 2) SDRX_TC_GO(9, 2); else SDRX_TC_GO(9, 1); }
#undef SDRX_TC_GO
    if (ev1) CU(cudaEventRecord(ev1, st));
    CU(cudaEventRecord(tc.ev_b_free[par], st));
    tc.b_busy[par] = true;
    *launches += 1;
    CU(cudaGetLastError());
    if (tc.d_trace && tc.trace_left > 0 && --tc.trace_left == 0) {       // debug only: the 40th launch's role timings to stderr
        CU(cudaStreamSynchronize(st));
        const int n_cta = (int)(grid.x * grid.y * grid.z);
        std::vector<unsigned long long> h((size_t)n_cta * 24);
        CU(cudaMemcpy(h.data(), tc.d_trace, h.size() * 8, cudaMemcpyDeviceToHost));
        const char *names[3] = {"epilogue warp 0", "TMA producer", "MMA issuer"};
        for (int r = 0; r < 3; ++r) {
            double tot = 0, w = 0, w2 = 0, s1 = 0, s2 = 0;
            for (int c = 0; c < n_cta; ++c) {
                const unsigned long long *o = &h[((size_t)c * 3 + r) * 8];
                tot += (double)o[0]; w += (double)o[1]; w2 += (double)o[2]; s1 += (double)o[3]; s2 += (double)o[4];
            }
            fprintf(stderr, "[tc trace] %-16s cycles %9.0f  waiting %9.0f (%s)  waiting %9.0f (%s)  mma issue %9.0f  commits %9.0f\n", names[r], tot / n_cta,
                    w / n_cta, r == 0 ? "accumulator full" : r == 1 ? "stage empty" : "stage full", w2 / n_cta, r == 2 ? "accumulator empty" : "-",
                    s1 / n_cta, s2 / n_cta);
        }
    }
    return SDRX_OK;
}

}  // namespace sdrx

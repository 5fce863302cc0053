// sdrx_b200 -- the C ABI (include/sdrx_b200.h) over the CUDA kernels.  Host side only orchestrates:
// pinned double-buffered ingest, stream/event plumbing, block-to-block carries, state save/load.
// There is no CPU compute path: every sample is touched by a kernel in bank_*.cuh / demod.cuh.
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "host_common.h"
#include "bank_generic.cuh"
#include "bank_fast.cuh"
#include "demod.cuh"
#include "polyphase.cuh"
#include "tgemm.cuh"
#include "tgemm_tc.cuh"

using namespace sdrx;
static_assert(kFastSideSms == kSideSms, "the bank leaves the SMs the side kernels expect");

namespace {

constexpr int kNBuf = 3;          // raw input ring: block b lands in d_raw[b % 3] while block b-1's prep / fold and block b-2's readers still run
constexpr int kNOut = 10;         // ring of output sets: a streaming caller may run up to kNOut-1 blocks ahead of the results it reads
constexpr int kPwrPartials = 296;     // power CTAs of k_prep: two per SM
constexpr uint32_t kStateMagic = 0x58524453u;  // "SDRX"

struct OutSet {                   // device-side results of one block
    float2  *ch_iq = nullptr;     // [C][512]
    float   *audio = nullptr;     // [512][Cpad] SAMPLE-major (k_chan_lfagc's lanes are channels); rows leave through k_export_audio
    float   *snr = nullptr;       // [C]
    float   *mix_pre = nullptr, *mix = nullptr;   // [1024]
    int16_t *pcm = nullptr;       // [1024]
    float   *pwr = nullptr;       // [1]
    uint8_t *sql = nullptr;       // [C] snapshot of the state after the decision
    float   *agc_gain = nullptr, *lf_gain = nullptr;   // [C] snapshots
    uint8_t *mixed = nullptr;     // [Cpad] 1 where this block's audio goes into the mix (k_chan_recur -> k_mix_fir2_pcm)
    CUtensorMap iq_map;           // ch_iq as the AGC's TMA source: FLOAT32 [C][2 * 512], 32 x 160 boxes, 128 B swizzle
    int     *audio_idx = nullptr; // [C] channels whose audio is mixed this block, ascending (k_mix_fir2_pcm)
    int     *n_audio = nullptr;   // [1] how many
};

}  // namespace

struct sdrx_engine {
    int dev = 0;
    unsigned flags = 0;
    int C = 0, Cpad = 0, N = 0, S = 0;   // C: channels over all streams (NS * Cs)
    int NS = 1, Cs = 0;                  // streams, channels per stream (sdrx_engine_config.n_streams / n_channels)
    size_t raw_stream_bytes = 0;         // one stream's [raw_hist | block_len] slot of a raw buffer
    int m[SDRX_MAX_STAGES] = {0}, H[SDRX_MAX_STAGES] = {0}, tap_off[SDRX_MAX_STAGES] = {0};
    std::vector<float> taps;
    unsigned block_len = 0;
    int raw_hist = 0;                    // history samples kept in front of each raw block
    long stage_stride[SDRX_MAX_STAGES] = {0};   // per-channel stride of stage s input buffer (s >= 1)
    long chiq_stride = kChunk;           // per-channel stride of the chain output
    AgcParams agc{}, lf{};
    float agc_power0 = 0.25f, lf_power0 = 0.0625f;
    int lf_active = 0;
    int fir2_ntaps = 0;

    cudaStream_t s_main = nullptr, s_copy = nullptr;
    cudaStream_t s_prep = nullptr;        // k_prep (+ the tensor-core fold) of a block: the next block's H2D runs beside them on s_copy
    cudaEvent_t ev_h2d[kNBuf] = {}, ev_prep[kNBuf] = {};   // block landed in d_raw[b]; its delay-line prefix is in place
    bool prep_pending[kNBuf] = {};
    // Boundary B runs under the following blocks' channel banks as a three-deep pipeline, one stream per kernel: the
    // detector of block b+2, the recurrences of block b+1 and the mix of block b overlap (each kernel only depends on
    // its own predecessor -- squelch state, AGC state, FIR2 ring -- and on the previous kernel of its own block), so a
    // step costs the longest of them, not their sum
    cudaStream_t s_side = nullptr;        // k_chan_head<kHeadDetect>
    cudaStream_t s_recur = nullptr;       // k_chan_agc
    cudaStream_t s_lf = nullptr;          // k_chan_env + k_chan_lfagc: block b's LfAGC runs next to block b + 1's AGC
    cudaStream_t s_mix = nullptr;         // k_mix_fir2_pcm
    cudaEvent_t ev_head_done = nullptr;   // the last MSD stage of the current block has been queued behind this point (main -> side)
    cudaEvent_t ev_det_done[kNOut] = {}, ev_recur_done[kNOut] = {};
    cudaEvent_t ev_agc_done[2] = {}, ev_env_done[2] = {};   // d_agc_iq[par] written / read (ping-pong between the two recurrence streams)
    bool env_pending[2] = {false, false};
    unsigned agc_par = 0;
    cudaEvent_t ev_side_done[kNOut] = {};   // all of boundary B that reads / writes output set b is complete (recorded behind the mix)
    bool side_pending[kNOut] = {};
    cudaStream_t s_out = nullptr;         // asynchronous result fetch (sdrx_engine_fetch_outputs): D2H overlaps the next block
    cudaEvent_t ev_fetch_done[kNOut] = {};
    bool fetch_pending[kNOut] = {};
    uint8_t *d_gate[kNOut] = {};            // [C] squelch state entering the block (k_chan_head -> k_chan_recur)
    cudaEvent_t ev_copied[kNBuf] = {};
    cudaEvent_t ev_raw_free[kNBuf] = {};   // last reader of d_raw[b] has been queued
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr, ev_t2 = nullptr, ev_b0 = nullptr, ev_b1 = nullptr;
    bool bank_timed = false;
    bool timing = false, timing_valid = false;

    // device memory
    uint8_t *d_raw[kNBuf] = {};
    size_t raw_buf_bytes = 0;
    float2 *d_table_t = nullptr;         // [N][Cpad]
    uint8_t *d_has_tr = nullptr;
    float *d_taps = nullptr;
    float2 *d_stage_in[SDRX_MAX_STAGES] = {nullptr};
    float *d_pwr_partial[kNOut] = {};      // per parity: k_prep of block b+1 runs while block b's mix kernel reads its own
    OutSet out[kNOut];
    // boundary B
    ChannelState st{};
    DemodTables tb{};
    float *d_window = nullptr, *d_shape = nullptr, *d_ramp_up = nullptr, *d_ramp_down = nullptr;
    float2 *d_twiddle = nullptr;
    uint16_t *d_bitrev = nullptr;
    float *d_sql_level = nullptr;
    int *d_pos = nullptr;
    uint8_t *d_mod = nullptr;             // [C] Channel::mod (sdrx.cpp:230)
    float *d_c_adj = nullptr, *d_fir_hist = nullptr;
    float *d_parts = nullptr;             // staging of the per-GPU partial mixes (sdrx_engine_finish_mix)
    size_t parts_cap = 0;
    float2 *d_agc_iq[2] = {nullptr, nullptr};   // [512][Cpad] gain-adjusted channel IQ, sample-major (k_chan_agc -> k_chan_env), ping-pong
    float *d_env[2] = {nullptr, nullptr}; // [512][Cpad] envelope / discriminator output (k_chan_env -> k_chan_lfagc), ping-pong
    float2 *d_fm[2] = {nullptr, nullptr}; // [C] Demod::prev_sample (sdrx.cpp:202) as a ping-pong pair: d_fm[0] is st.fm_prev's allocation; block b's
                                          // envelope reads d_fm[fm_cur] and writes the other one for every channel
    int fm_cur = 0;
    // debug timeline (sdrx_debug_timeline): completion time of every kernel of the last kTlBlocks blocks
    static constexpr int kTlBlocks = 16, kTlKinds = 10;
    bool tl_on = false;
    cudaEvent_t tl_ev[kTlBlocks][kTlKinds] = {};
    cudaEvent_t tl_base = nullptr;
    uint64_t tl_first = 0;                // e->blocks when the timeline was switched on
    cudaEvent_t ev_ext = nullptr;         // sdrx_engine_wait_stream: marks the caller's stream
    static constexpr int kNFin = 4;       // sdrx_engine_finish_mix_async results in flight
    float *d_fin_mix_pre[kNFin] = {}, *d_fin_mix[kNFin] = {};
    int16_t *d_fin_pcm[kNFin] = {};
    uint64_t n_fin = 0;
    unsigned *d_ticket = nullptr;         // arrival counter of k_mix_fir2_pcm's CTAs (the last one runs FIR2); self-resetting
    float *d_rows_scratch[2] = {nullptr, nullptr};   // [C][512] audio rows on their way to a dense / pageable host array: one per
                                                     // stream that copies results out (blocking gets: main, fetches: s_out)
    FastBank fast;                        // specialised channel bank (bank_fast.cuh)
    bool use_fast = false;
    TableGemm tg;                         // fold + table GEMM (tgemm.cuh): SDRX_FLAG_TGEMM
    bool use_tgemm = false;
    TcGemm tc;                            // ... its GEMM on the tensor cores (tgemm_tc.cuh): SDRX_FLAG_TGEMM_TC
    bool use_tc = false;
    int mat_len = 625, mat_stages = 2;    // materialize_generic: raw samples it replays, leading stages it runs
    bool tail_fusable = false;            // last stage as one per-channel kernel fed by the bank's rows + tail fix-ups
    bool tail_fast = false;               // ... in its specialised form (67 taps, decimate by 5: k_chan_fir_fast)
    FirTaps<67> tail_taps{};
    // The specialised last stage (k_chan_fir_fast) runs on the detector's stream, UNDER the next block's bank (the bank's
    // staging ring leaves room for two of its CTAs per SM): the bank of block b+1 writes one stage-row / tails buffer while
    // the last stage of block b reads the other and slides its delay line into the partner.
    float2 *d_stage_alt = nullptr;        // ping-pong partner of d_stage_in[S-1]
    cudaEvent_t ev_tail_done[2] = {nullptr, nullptr};   // last-stage kernel that read row / tails buffer `par` is complete
    bool tail_pending[2] = {false, false};
    int tail_par = 0;                     // row / tails buffer the next fused block uses
    int last_tail_par = -1;               // buffer of the most recent side-stream last stage (-1: none outstanding)
    bool poly = false;                    // SDRX_FLAG_POLYPHASE: boundary A is k_poly_channelizer
    int poly_D = 0, poly_Lh = 0;          // total decimation, composite prototype length
    float *d_hc = nullptr;                // [poly_Lh] h1(z) h2(z^5) h3(z^25) h4(z^125)
    float2 *d_wtab = nullptr;             // [1200] exp(-2 pi i n / 1200)
    int *d_bin = nullptr;                 // [C] DFT bin of every channel: ch_offset mod 1200
    bool ftfir = false;                   // --ftfir: stage 0 is the frequency-translating FIR (msd.hpp:104-131)
    int ft_sets = 0;                      // coefficient sets, N / m[0] (msd.hpp:196)
    float2 *d_hk = nullptr;               // [ft_sets * H0][Cpad]

    // pinned staging for pageable callers
    uint8_t *h_stage[kNBuf] = {};
    bool stage_busy[kNBuf] = {};

    // carried host-side state
    long trans_pos = 0;                  // MSD::trans_pos_ (msd.hpp:440), same for every channel
    int phase[SDRX_MAX_STAGES] = {0};    // inputs seen since the stage's last output (m - isn_)
    int ft_k = 0;                        // S::k_ of stage 0 (msd.hpp:434), same for every channel
    int fmt = -1;                        // wire format of the stream so far
    size_t agc_smem = kAgcSmem, lf_smem = kLfSmem;   // dynamic shared memory the recurrence kernels are launched with
    int agc_bufs = 3, lf_bufs = 3;                   // ... and their tile ring depths
    int ring = kNOut;                    // output sets in use (<= kNOut; SDRX_RING overrides it for experiments)
    uint64_t blocks = 0;                 // output sets handed out so far (push_block and demod_block): selects out[blocks % e->ring]
    uint64_t pushes = 0;                 // blocks through the channelizer: its parity selects the raw ping-pong buffer
    unsigned prev_n = 0;                 // samples of the previous block (for the raw carry)
    uint64_t total = 0;                  // samples pushed so far (what of the raw history belongs to the stream)
    bool generic_stale = false;          // stage 1-2 delay lines not maintained (the specialised bank ran last)
    unsigned last_out_len = 0;
    bool last_demod = false;
    uint64_t launches = 0;
};

// ------------------------------------------------------------------------------------------------
namespace sdrx {
std::string &last_error_ref() {
    static thread_local std::string err;
    return err;
}
}  // namespace sdrx

extern "C" const char *sdrx_last_error(void) { return last_error_ref().c_str(); }
extern "C" const char *sdrx_version(void) { return "sdrx_b200 0.1 (sm_100a)"; }
extern "C" int sdrx_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

// sdrx.cpp:1682-1690: the expression is evaluated in double (M_PI), narrowed to float by the
// std::complex<float> constructor, then std::exp(std::complex<float>) of this host's libm.
extern "C" void sdrx_make_translator(int N, int z, int ch_offset, float *out) {
    if (ch_offset == 0) return;
    for (int n = 0; n < N; n++) {
        std::complex<float> e(0.0f, -2.0f * M_PI * n * ch_offset * (float)z / (float)N);
        std::complex<float> v = std::exp(e);
        out[2 * n] = v.real();
        out[2 * n + 1] = v.imag();
    }
}

// sdrx.cpp:898-900
extern "C" void sdrx_make_window(float *out) {
    for (unsigned n = 0; n < (unsigned)kChunk; n++) out[n] = 0.54f - 0.46f * std::cos((2.0f * M_PI * n) / kChunk);
}

extern "C" void *sdrx_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes) != cudaSuccess) { last_error_ref() = "cudaMallocHost failed"; cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void sdrx_host_free(void *p) { if (p) cudaFreeHost(p); }
extern "C" void *sdrx_device_alloc(int device, size_t bytes) {
    DevGuard g(device);
    void *p = nullptr;
    if (!g.ok || cudaMalloc(&p, bytes) != cudaSuccess) { last_error_ref() = "cudaMalloc failed"; cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void sdrx_device_free(int device, void *p) { DevGuard g(device); if (p) cudaFree(p); }
extern "C" int sdrx_device_upload(int device, void *d_dst, const void *h_src, size_t bytes) {
    DevGuard g(device);
    CU(cudaMemcpy(d_dst, h_src, bytes, cudaMemcpyHostToDevice));
    return SDRX_OK;
}

// ------------------------------------------------------------------------------------------------
namespace {

// The tensor map k_chan_agc reads a block's channel IQ through (demod.cuh).  The encoder lives in the driver library: it
// is fetched through the runtime (cudaGetDriverEntryPoint), so nothing links against libcuda.
int make_map_2d(CUtensorMap *map, void *base, cuuint64_t inner, cuuint64_t rows, cuuint64_t pitch_bytes, cuuint32_t box_inner, cuuint32_t box_rows,
                CUtensorMapDataType dtype = CU_TENSOR_MAP_DATA_TYPE_FLOAT32) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn encode = nullptr;
    if (!encode) {
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult q;
        CU(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
        if (!fn || q != cudaDriverEntryPointSuccess) return fail(SDRX_ERR_CUDA, "cuTensorMapEncodeTiled is not available in this driver");
        encode = (encode_fn)fn;
    }
    const cuuint64_t dims[2] = {inner, rows};
    const cuuint64_t strides[1] = {pitch_bytes};
    const cuuint32_t box[2] = {box_inner, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(map, dtype, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(SDRX_ERR_CUDA, "cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + ")");
    return SDRX_OK;
}

template <typename T>
int dalloc(T **p, size_t count) {
    CU(cudaMalloc((void **)p, std::max<size_t>(count, 1) * sizeof(T)));
    CU(cudaMemset(*p, 0, std::max<size_t>(count, 1) * sizeof(T)));
    return SDRX_OK;
}
template <typename T>
int upload(T *d, const T *h, size_t count) {
    CU(cudaMemcpy(d, h, count * sizeof(T), cudaMemcpyHostToDevice));
    return SDRX_OK;
}

void agc_defaults(AgcParams &p, float attack, float decay, float reference, float max_gain, float alpha,
                  float d_max_gain, float d_alpha) {
    // zero fields take what sdrx.cpp:1702-1710 sets (reference 1, attack 1, decay 0.01)
    p.attack = attack != 0.0f ? attack : 1.0f;
    p.decay = decay != 0.0f ? decay : 0.01f;
    p.reference = reference != 0.0f ? reference : 1.0f;
    p.max_gain = max_gain != 0.0f ? max_gain : d_max_gain;
    p.alpha = alpha != 0.0f ? alpha : d_alpha;
    p.min_gain = (p.max_gain > 0.0f) ? 1.0f / p.max_gain : 1.0f / 200.0f;   // agc.hpp:34-35, :41-44
}

int reset_state(sdrx_engine *e) {
    const size_t C = (size_t)e->C;
    CU(cudaStreamSynchronize(e->s_side));
    CU(cudaStreamSynchronize(e->s_recur));
    CU(cudaStreamSynchronize(e->s_lf));
    CU(cudaStreamSynchronize(e->s_mix));
    CU(cudaStreamSynchronize(e->s_out));
    for (int b = 0; b < kNOut; ++b) e->side_pending[b] = e->fetch_pending[b] = false;
    e->tail_pending[0] = e->tail_pending[1] = false;
    e->tail_par = 0; e->last_tail_par = -1;
    if (e->d_stage_alt) CU(cudaMemsetAsync(e->d_stage_alt, 0, sizeof(float2) * (size_t)e->stage_stride[e->S - 1] * C, e->s_main));
    for (int b = 0; b < kNBuf; ++b) CU(cudaMemsetAsync(e->d_raw[b], 0, e->raw_buf_bytes, e->s_main));
    for (int s = 1; s < e->S; ++s)
        if (e->d_stage_in[s]) CU(cudaMemsetAsync(e->d_stage_in[s], 0, sizeof(float2) * (size_t)e->stage_stride[s] * C, e->s_main));
    if (!(e->flags & SDRX_FLAG_NO_DEMOD)) {
        std::vector<float> ones(C, 1.0f), p0(C, e->agc_power0), p1(C, e->lf_power0);
        CU(cudaMemcpyAsync(e->st.agc_gain, ones.data(), C * 4, cudaMemcpyHostToDevice, e->s_main));
        CU(cudaMemcpyAsync(e->st.lf_gain, ones.data(), C * 4, cudaMemcpyHostToDevice, e->s_main));
        CU(cudaMemcpyAsync(e->st.agc_power, p0.data(), C * 4, cudaMemcpyHostToDevice, e->s_main));
        CU(cudaMemcpyAsync(e->st.lf_power, p1.data(), C * 4, cudaMemcpyHostToDevice, e->s_main));
        CU(cudaMemsetAsync(e->st.sql_state, 0, C, e->s_main));
        CU(cudaMemsetAsync(e->st.sql_prev, 0, C, e->s_main));
        for (int b = 0; b < kNOut; ++b) CU(cudaMemsetAsync(e->out[b].mixed, 0, (size_t)e->Cpad, e->s_main));
        CU(cudaMemsetAsync(e->d_fm[0], 0, C * sizeof(float2), e->s_main));
        CU(cudaMemsetAsync(e->d_fm[1], 0, C * sizeof(float2), e->s_main));
        CU(cudaMemsetAsync(e->d_fir_hist, 0, sizeof(float) * e->NS * 2 * (size_t)std::max(e->fir2_ntaps - 1, 1), e->s_main));
        CU(cudaMemsetAsync(e->d_ticket, 0, 2 * sizeof(unsigned) * e->NS, e->s_main));
        CU(cudaStreamSynchronize(e->s_main));   // the host vectors above die here
    }
    CU(cudaStreamSynchronize(e->s_main));
    e->trans_pos = 0;
    e->ft_k = 0;
    for (int s = 0; s < SDRX_MAX_STAGES; ++s) e->phase[s] = 0;
    e->fmt = -1;
    e->blocks = 0;
    e->pushes = 0;
    e->prev_n = 0;
    e->total = 0;
    e->generic_stale = false;
    e->last_out_len = 0;
    e->last_demod = false;
    e->timing_valid = false;
    return SDRX_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
extern "C" void sdrx_engine_destroy(sdrx_engine *e) {
    if (!e) return;
    DevGuard g(e->dev);
    cudaDeviceSynchronize();
    for (int b = 0; b < kNBuf; ++b) {
        cudaFree(e->d_raw[b]);
        if (e->h_stage[b]) cudaFreeHost(e->h_stage[b]);
        if (e->ev_copied[b]) cudaEventDestroy(e->ev_copied[b]);
        if (e->ev_raw_free[b]) cudaEventDestroy(e->ev_raw_free[b]);
    }
    for (int b = 0; b < kNOut; ++b) {
        OutSet &o = e->out[b];
        cudaFree(o.ch_iq); cudaFree(o.audio); cudaFree(o.snr); cudaFree(o.mix_pre); cudaFree(o.mix);
        cudaFree(o.pcm); cudaFree(o.pwr); cudaFree(o.sql); cudaFree(o.agc_gain); cudaFree(o.lf_gain);
        cudaFree(o.audio_idx); cudaFree(o.n_audio); cudaFree(o.mixed);
        if (e->ev_det_done[b]) cudaEventDestroy(e->ev_det_done[b]);
        if (e->ev_recur_done[b]) cudaEventDestroy(e->ev_recur_done[b]);
        if (e->ev_side_done[b]) cudaEventDestroy(e->ev_side_done[b]);
        if (e->ev_fetch_done[b]) cudaEventDestroy(e->ev_fetch_done[b]);
        cudaFree(e->d_gate[b]); cudaFree(e->d_pwr_partial[b]);
    }
    if (e->ev_head_done) cudaEventDestroy(e->ev_head_done);
    for (int b = 0; b < 2; ++b) if (e->ev_tail_done[b]) cudaEventDestroy(e->ev_tail_done[b]);
    cudaFree(e->d_stage_alt);
    for (int s = 0; s < SDRX_MAX_STAGES; ++s) cudaFree(e->d_stage_in[s]);
    cudaFree(e->d_table_t); cudaFree(e->d_has_tr); cudaFree(e->d_taps); cudaFree(e->d_hk);
    cudaFree(e->st.agc_gain); cudaFree(e->st.agc_power); cudaFree(e->st.lf_gain); cudaFree(e->st.lf_power);
    cudaFree(e->d_fm[0]); cudaFree(e->d_fm[1]); cudaFree(e->d_mod);
    cudaFree(e->st.sql_state); cudaFree(e->st.sql_prev);
    cudaFree(e->d_window); cudaFree(e->d_shape); cudaFree(e->d_ramp_up); cudaFree(e->d_ramp_down);
    cudaFree(e->d_twiddle); cudaFree(e->d_bitrev); cudaFree(e->d_sql_level); cudaFree(e->d_pos);
    cudaFree(e->d_c_adj); cudaFree(e->d_fir_hist); cudaFree(e->d_parts); cudaFree(e->d_ticket); cudaFree(e->d_rows_scratch[0]); cudaFree(e->d_rows_scratch[1]);
    cudaFree(e->d_agc_iq[0]); cudaFree(e->d_agc_iq[1]); cudaFree(e->d_env[0]); cudaFree(e->d_env[1]);
    for (int b = 0; b < 2; ++b) { if (e->ev_agc_done[b]) cudaEventDestroy(e->ev_agc_done[b]); if (e->ev_env_done[b]) cudaEventDestroy(e->ev_env_done[b]); }
    cudaFree(e->d_hc); cudaFree(e->d_wtab); cudaFree(e->d_bin);
    for (int b = 0; b < sdrx_engine::kNFin; ++b) { cudaFree(e->d_fin_mix_pre[b]); cudaFree(e->d_fin_mix[b]); cudaFree(e->d_fin_pcm[b]); }
    if (e->ev_ext) cudaEventDestroy(e->ev_ext);
    if (e->tl_base) cudaEventDestroy(e->tl_base);
    for (auto &row : e->tl_ev) for (cudaEvent_t ev : row) if (ev) cudaEventDestroy(ev);
    e->fast.destroy();
    e->tg.destroy();
    e->tc.destroy();
    if (e->ev_t0) cudaEventDestroy(e->ev_t0);
    if (e->ev_t1) cudaEventDestroy(e->ev_t1);
    if (e->ev_t2) cudaEventDestroy(e->ev_t2);
    if (e->ev_b0) cudaEventDestroy(e->ev_b0);
    if (e->ev_b1) cudaEventDestroy(e->ev_b1);
    if (e->s_main) cudaStreamDestroy(e->s_main);
    if (e->s_copy) cudaStreamDestroy(e->s_copy);
    if (e->s_prep) cudaStreamDestroy(e->s_prep);
    for (int b = 0; b < kNBuf; ++b) { if (e->ev_h2d[b]) cudaEventDestroy(e->ev_h2d[b]); if (e->ev_prep[b]) cudaEventDestroy(e->ev_prep[b]); }
    if (e->s_side) cudaStreamDestroy(e->s_side);
    if (e->s_recur) cudaStreamDestroy(e->s_recur);
    if (e->s_lf) cudaStreamDestroy(e->s_lf);
    if (e->s_mix) cudaStreamDestroy(e->s_mix);
    if (e->s_out) cudaStreamDestroy(e->s_out);
    delete e;
}

namespace {

int create_impl(const sdrx_engine_config *cfg, sdrx_engine *e) {
    const int NS = std::max(cfg->n_streams, 1), Cs = cfg->n_channels;
    const int C = NS * Cs, S = cfg->n_stages;            // per-channel arrays below: stream-major, the plan repeated per stream
    e->dev = cfg->device;
    if (const char *er = getenv("SDRX_RING")) e->ring = std::min(kNOut, std::max(3, atoi(er)));
    e->flags = cfg->flags;
    e->NS = NS; e->Cs = Cs;
    e->C = C;
    e->Cpad = (C + 31) / 32 * 32;
    e->N = std::max(cfg->translator_len, 1);
    e->S = S;
    e->block_len = cfg->block_len;
    int off = 0;
    for (int s = 0; s < S; ++s) {
        e->m[s] = cfg->m[s];
        e->H[s] = cfg->ntaps[s];
        e->tap_off[s] = off;
        off += cfg->ntaps[s];
    }
    e->taps.assign(cfg->taps, cfg->taps + off);
    e->raw_hist = std::max((e->H[0] - 1 + 7) / 8 * 8, 128);
    const bool eligible = fast_eligible(S, e->m, e->H, e->N);
    e->ftfir = (cfg->flags & SDRX_FLAG_FTFIR) && cfg->translators && cfg->translator_len > 0;
    e->poly = (cfg->flags & SDRX_FLAG_POLYPHASE) != 0;
    if (e->poly) {
        if (!cfg->ch_offset) return fail(SDRX_ERR_ARG, "SDRX_FLAG_POLYPHASE needs ch_offset");
        if (e->N != kPolyN || e->ftfir || NS > 1 || (cfg->flags & SDRX_FLAG_NO_DEMOD))
            return fail(SDRX_ERR_UNSUPPORTED, "SDRX_FLAG_POLYPHASE: 10 MS/s chain (translator period 1200), one stream, with demodulation, no --ftfir");
        // the four stages as one prototype, in double: h_c(z) = prod_s h_s(z^(m_0 .. m_(s-1)))
        std::vector<double> hc(1, 1.0);
        long up = 1;
        for (int st = 0; st < S; ++st) {
            std::vector<double> nx((hc.size() - 1) + (size_t)(e->H[st] - 1) * up + 1, 0.0);
            for (size_t i = 0; i < hc.size(); ++i)
                for (int j = 0; j < e->H[st]; ++j) nx[i + (size_t)j * up] += hc[i] * (double)e->taps[e->tap_off[st] + j];
            hc.swap(nx);
            up *= e->m[st];
        }
        e->poly_D = (int)up;
        e->poly_Lh = (int)hc.size();
        std::vector<float> hcf(hc.begin(), hc.end());
        std::vector<float2> wt(kPolyN);
        for (int n = 0; n < kPolyN; ++n) {
            const double a = -2.0 * M_PI * (double)n / (double)kPolyN;
            wt[n] = make_float2((float)cos(a), (float)sin(a));
        }
        std::vector<int> bin(C);
        for (int c = 0; c < C; ++c) bin[c] = ((cfg->ch_offset[c] % kPolyN) + kPolyN) % kPolyN;
        int rc;
        if ((rc = dalloc(&e->d_hc, hcf.size())) || (rc = upload(e->d_hc, hcf.data(), hcf.size())) ||
            (rc = dalloc(&e->d_wtab, wt.size())) || (rc = upload(e->d_wtab, wt.data(), wt.size())) ||
            (rc = dalloc(&e->d_bin, bin.size())) || (rc = upload(e->d_bin, bin.data(), bin.size())))
            return rc;
        // output M of a block needs the poly_Lh - 1 samples before sample D M + D - 1: poly_Lh - D of them before the block
        e->raw_hist = std::max(e->raw_hist, (e->poly_Lh - e->poly_D + 7) / 8 * 8);
        if ((rc = raise_dyn_smem((const void *)k_poly_channelizer<SDRX_FMT_U8>, poly_smem_bytes())) ||
            (rc = raise_dyn_smem((const void *)k_poly_channelizer<SDRX_FMT_S16>, poly_smem_bytes())) ||
            (rc = raise_dyn_smem((const void *)k_poly_channelizer<SDRX_FMT_F32>, poly_smem_bytes())))
            return rc;
    }
    const bool want_fast = !(cfg->flags & SDRX_FLAG_FORCE_GENERIC) && eligible && !e->ftfir && !e->poly;
    if (eligible) e->raw_hist = std::max(e->raw_hist, kFastRawHist);   // also when forced generic: state blobs stay portable
    // fold + table GEMM (opt-in): any chain with translator tables of a period that is a multiple of 16 (1200 at 10 / 6 MS/s,
    // 288 at 2.4 MS/s ...)
    const bool want_tgemm = (cfg->flags & (SDRX_FLAG_TGEMM | SDRX_FLAG_TGEMM_TC)) && !(cfg->flags & (SDRX_FLAG_FORCE_GENERIC | SDRX_FLAG_POLYPHASE)) && !e->ftfir && S >= 2 &&
                            cfg->translators && cfg->translator_len >= kGemmK && cfg->translator_len % kGemmK == 0;
    if (want_tgemm) {
        long d_all = 1, d_head = 1;
        for (int s = 0; s < S; ++s) { d_all *= e->m[s]; if (s < S - 1) d_head *= e->m[s]; }
        e->mat_len = (int)(d_head * (e->H[S - 1] + 7));          // rebuilds every stage's delay line from raw samples
        e->mat_stages = S - 1;
        int lh = 1; long up = 1;
        for (int s = 0; s < S; ++s) { lh += (e->H[s] - 1) * (int)up; up *= e->m[s]; }
        e->raw_hist = std::max(e->raw_hist, (std::max(lh - (int)d_all, e->mat_len) + 7) / 8 * 8);
    }

    {
        // experiment knob: SDRX_MAIN_PRIO = 1 puts the channelizer stream at the highest priority too (its whole-SM CTAs first in line)
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        const char *mp = getenv("SDRX_MAIN_PRIO");
        CU(cudaStreamCreateWithPriority(&e->s_main, cudaStreamNonBlocking, (mp && atoi(mp) == 1) ? hi : lo));
    }
    CU(cudaStreamCreateWithFlags(&e->s_copy, cudaStreamNonBlocking));
    {
        // k_prep and the tensor-core fold are short, latency-critical links of the block-to-block loop (the next GEMM waits for the
        // fold, the next k_prep for this fold's stream slot): at the default priority their CTAs queued behind the detector's 960 and
        // took 22 + 40 us in the pipeline for 6 + 17 us of work -- and that loop, not boundary B, set the step
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        const char *pp = getenv("SDRX_PREP_PRIO");
        CU(cudaStreamCreateWithPriority(&e->s_prep, cudaStreamNonBlocking, (pp && atoi(pp) == 0) ? lo : hi));
    }
    for (int b = 0; b < kNBuf; ++b) {
        CU(cudaEventCreateWithFlags(&e->ev_h2d[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_prep[b], cudaEventDisableTiming));
    }
    CU(cudaStreamCreateWithFlags(&e->s_side, cudaStreamNonBlocking));
    {
        // the recurrences and the mix are few, long CTAs that need most of a free SM's shared memory; the detector is 960
        // short ones that would otherwise keep nibbling at it: higher priority puts the long ones first in line
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&e->s_recur, cudaStreamNonBlocking, hi));
        CU(cudaStreamCreateWithPriority(&e->s_lf, cudaStreamNonBlocking, hi));
        CU(cudaStreamCreateWithPriority(&e->s_mix, cudaStreamNonBlocking, hi));
    }
    CU(cudaStreamCreateWithFlags(&e->s_out, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&e->ev_head_done, cudaEventDisableTiming));
    for (int b = 0; b < 2; ++b) {
        CU(cudaEventCreateWithFlags(&e->ev_agc_done[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_env_done[b], cudaEventDisableTiming));
    }
    CU(cudaEventCreateWithFlags(&e->ev_ext, cudaEventDisableTiming));
    for (int b = 0; b < 2; ++b) CU(cudaEventCreateWithFlags(&e->ev_tail_done[b], cudaEventDisableTiming));
    for (int b = 0; b < kNOut; ++b) {
        CU(cudaEventCreateWithFlags(&e->ev_side_done[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_det_done[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_recur_done[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_fetch_done[b], cudaEventDisableTiming));
    }
    for (int b = 0; b < kNBuf; ++b) {
        CU(cudaEventCreateWithFlags(&e->ev_copied[b], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&e->ev_raw_free[b], cudaEventDisableTiming));
    }
    CU(cudaEventCreate(&e->ev_t0));
    CU(cudaEventCreate(&e->ev_t1));
    CU(cudaEventCreate(&e->ev_t2));
    CU(cudaEventCreate(&e->ev_b0));
    CU(cudaEventCreate(&e->ev_b1));

    // raw ping-pong: [raw_hist | block_len] samples at the widest wire format
    e->raw_stream_bytes = (((size_t)e->raw_hist + e->block_len) * 8 + 255) / 256 * 256;
    e->raw_buf_bytes = e->raw_stream_bytes * NS + 64;
    for (int b = 0; b < kNBuf; ++b) {
        int rc = dalloc(&e->d_raw[b], e->raw_buf_bytes);
        if (rc) return rc;
    }

    // translator, transposed to [N][Cpad]; (1,0) rows are never read (has_tr == 0 skips the multiply)
    {
        std::vector<float2> t((size_t)e->N * e->Cpad, make_float2(1.0f, 0.0f));
        std::vector<uint8_t> has(e->Cpad, 0);
        for (int c = 0; c < C; ++c) {
            const int cs = c % Cs;
            const bool on = cfg->translators && cfg->translator_len > 0 &&
                            (cfg->has_translator ? cfg->has_translator[cs] != 0 : true);
            has[c] = on ? 1 : 0;
            if (!on) continue;
            const float *row = cfg->translators + (size_t)cs * cfg->translator_len * 2;
            for (int n = 0; n < e->N; ++n) t[(size_t)n * e->Cpad + c] = make_float2(row[2 * n], row[2 * n + 1]);
        }
        int rc = dalloc(&e->d_table_t, t.size());
        if (rc) return rc;
        if ((rc = upload(e->d_table_t, t.data(), t.size()))) return rc;
        if ((rc = dalloc(&e->d_has_tr, has.size()))) return rc;
        if ((rc = upload(e->d_has_tr, has.data(), has.size()))) return rc;
    }
    if (e->ftfir) {
        // msd.hpp:192-224: hk_[set][i] = h[i] * translator[(j + set*m + i) % N] * 2.0f  (float * complex, then
        // complex * float: every product rounds once, the doubling is exact), j = (N - (H-1) % N) % N
        const int N = e->N, H = e->H[0], m0 = e->m[0];
        e->ft_sets = N / m0;
        if (e->ft_sets < 1) return fail(SDRX_ERR_ARG, "ftfir: translator shorter than the first stage's decimation factor");
        std::vector<float2> hk((size_t)e->ft_sets * H * e->Cpad, make_float2(0.0f, 0.0f));
        const float *h = e->taps.data() + e->tap_off[0];
        for (int c = 0; c < C; ++c) {
            if (cfg->has_translator && !cfg->has_translator[c % Cs]) continue;
            const float *row = cfg->translators + (size_t)(c % Cs) * N * 2;
            size_t j = (size_t)((N - (H - 1) % N) % N);
            for (int set = 0; set < e->ft_sets; ++set) {
                size_t k = j;
                for (int i = 0; i < H; ++i) {
                    const float re = h[i] * row[2 * k], im = h[i] * row[2 * k + 1];
                    hk[((size_t)set * H + i) * e->Cpad + c] = make_float2(re * 2.0f, im * 2.0f);
                    if (++k == (size_t)N) k = 0;
                }
                j += (size_t)m0;
                if (j >= (size_t)N) j -= (size_t)N;
            }
        }
        int rc = dalloc(&e->d_hk, hk.size());
        if (rc) return rc;
        if ((rc = upload(e->d_hk, hk.data(), hk.size()))) return rc;
    }
    {
        int rc = dalloc(&e->d_taps, e->taps.size());
        if (rc) return rc;
        if ((rc = upload(e->d_taps, e->taps.data(), e->taps.size()))) return rc;
    }
    // stage input buffers for s >= 1: [C][H-1 + cap], cap = most inputs one call can deliver
    {
        size_t cap = e->block_len;
        for (int s = 1; s < S; ++s) {
            cap = cap / (size_t)e->m[s - 1] + 1;
            e->stage_stride[s] = (long)((e->H[s] - 1) + cap + 3) / 4 * 4;
            // a multi-stream engine only ever runs the specialised bank: it needs the last stage's rows, not the
            // generic kernels' stage-1 / stage-2 rows (4 GB at 64 x 128 channels)
            if (NS > 1 && s < S - 1) continue;
            int rc = dalloc(&e->d_stage_in[s], (size_t)e->stage_stride[s] * C);
            if (rc) return rc;
        }
    }
    {
        // the chain's output: 512 per channel for the reference's block sizes, more if a caller pushes
        // a longer block through the channelizer alone
        size_t cap = e->block_len;
        for (int s = 0; s < S; ++s) cap = cap / (size_t)e->m[s] + 1;
        e->chiq_stride = (long)std::max<size_t>(kChunk, (cap + 3) / 4 * 4);
        for (int b = 0; b < kNOut; ++b) {
            int rc = dalloc(&e->d_pwr_partial[b], (size_t)kPwrPartials * NS);
            if (rc) return rc;
            if ((rc = dalloc(&e->d_gate[b], C))) return rc;
        }
    }
    if (want_tgemm && (size_t)e->mat_len <= e->block_len) {
        std::vector<uint8_t> has_v((size_t)Cs, 0);
        for (int c = 0; c < Cs; ++c) has_v[c] = (cfg->has_translator ? cfg->has_translator[c] != 0 : true) ? 1 : 0;
        std::vector<int> partner = (cfg->flags & SDRX_FLAG_NO_PAIRING) ? std::vector<int>((size_t)Cs, -1)
                                                                       : fast_find_partners(cfg->translators, has_v.data(), Cs, e->N);
        size_t max_out = e->block_len;
        for (int s = 0; s < S; ++s) max_out /= (size_t)e->m[s];
        int rc = tgemm_init(e->tg, Cs, e->N, S, e->m, e->H, e->taps.data(), e->tap_off, cfg->translators, has_v.data(), partner,
                            (int)std::max<size_t>(max_out, 4), NS);
        if (rc) return rc;
        if ((rc = raise_dyn_smem((const void *)k_tgemm, kGemmSmem))) return rc;
        e->use_tgemm = true;
        if (cfg->flags & SDRX_FLAG_TGEMM_TC) {
            // experiment knobs (tools/diag_channelizers.py): plane products kept (9 exact / 6) and 16-entry k-steps per drain unit
            const char *ep = getenv("SDRX_TC_PRODUCTS"), *es = getenv("SDRX_TC_STEPS"), *ec = getenv("SDRX_TC_CLUSTER");
            auto bf16_map = [](CUtensorMap *map, void *base, uint64_t inner, uint64_t rows, uint64_t pitch, int box_inner, int box_rows) {
                return make_map_2d(map, base, inner, rows, pitch, (cuuint32_t)box_inner, (cuuint32_t)box_rows, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16);
            };
            if ((rc = tc_init(e->tc, e->tg, cfg->translators, has_v.data(), e->tg.h_slot_chan, ep ? atoi(ep) : 6, es ? atoi(es) : 1, ec ? atoi(ec) : 1, bf16_map)))
                return rc;
            if ((rc = raise_dyn_smem(tc_kernel_ptr(e->tc.products, e->tc.steps), kTcSmem))) return rc;
            if (const char *fd = getenv("SDRX_TC_FOLD_DOUBLE")) e->tc.fold_double = atoi(fd);
            if (const char *ed = getenv("SDRX_TC_DEBUG")) e->tc.debug = atoi(ed);   // timing experiments only: results are wrong
            if (e->tc.debug & 16) {
                CU(cudaMalloc((void **)&e->tc.d_trace, 8 * 24 * 4096));
                CU(cudaMemset(e->tc.d_trace, 0, 8 * 24 * 4096));
                e->tc.trace_left = 40;
            }
            e->use_tc = true;
        }
    } else if (want_tgemm) {
        e->mat_len = 625; e->mat_stages = 2;
    }
    if (want_fast && !e->use_tgemm) {
        int rc = fast_init(e->fast, Cs, e->N, e->taps.data() + e->tap_off[0], e->taps.data() + e->tap_off[1],
                           e->taps.data() + e->tap_off[2], cfg->translator_len > 0 ? cfg->translators : nullptr,
                           cfg->has_translator, !(cfg->flags & SDRX_FLAG_NO_PAIRING), NS);
        if (rc) return rc;
        e->use_fast = true;
    }
    for (int b = 0; b < kNOut; ++b) {
        OutSet &o = e->out[b];
        int rc;
        if ((rc = dalloc(&o.ch_iq, (size_t)C * e->chiq_stride))) return rc;
        if (e->flags & SDRX_FLAG_NO_DEMOD) continue;
        if ((rc = dalloc(&o.audio, (size_t)e->Cpad * kChunk))) return rc;
        if ((rc = make_map_2d(&o.iq_map, o.ch_iq, 2 * kChunk, (cuuint64_t)C, (cuuint64_t)e->chiq_stride * sizeof(float2), 2 * kRecurTile,
                              kAgcBoxRows)))
            return rc;
        if ((rc = dalloc(&o.snr, C))) return rc;
        if ((rc = dalloc(&o.mix_pre, (size_t)NS * 2 * kChunk))) return rc;
        if ((rc = dalloc(&o.mix, (size_t)NS * 2 * kChunk))) return rc;
        if ((rc = dalloc(&o.pcm, (size_t)NS * 2 * kChunk))) return rc;
        if ((rc = dalloc(&o.pwr, NS))) return rc;
        if ((rc = dalloc(&o.sql, C))) return rc;
        if ((rc = dalloc(&o.agc_gain, C))) return rc;
        if ((rc = dalloc(&o.lf_gain, C))) return rc;
        if ((rc = dalloc(&o.mixed, (size_t)e->Cpad))) return rc;
        if ((rc = dalloc(&o.audio_idx, C))) return rc;
        if ((rc = dalloc(&o.n_audio, NS))) return rc;
    }

    if (!(e->flags & SDRX_FLAG_NO_DEMOD)) {
        if (!cfg->fir2_taps || cfg->fir2_ntaps < 1 || !cfg->passband_shape || !cfg->ramp_up || !cfg->ramp_down)
            return fail(SDRX_ERR_ARG, "demod tables (fir2_taps, passband_shape, ramp_up, ramp_down) are required");
        if (cfg->fir2_ntaps > 2048) return fail(SDRX_ERR_UNSUPPORTED, "fir2_ntaps > 2048");
        agc_defaults(e->agc, cfg->agc_attack, cfg->agc_decay, cfg->agc_reference, cfg->agc_max_gain, cfg->agc_alpha,
                     300.0f, 0.01f);
        agc_defaults(e->lf, cfg->lf_attack, cfg->lf_decay, cfg->lf_reference, cfg->lf_max_gain, cfg->lf_alpha,
                     200.0f, 0.005f);
        e->agc_power0 = 0.25f;              // AGC(): power_(reference_) with the ctor default, agc.hpp:30-32
        e->lf_power0 = 0.25f * 0.25f;       // LfAGC(): power_(reference_*reference_), agc.hpp:86-89
        e->lf_active = cfg->use_lf_agc ? 1 : 0;
        e->fir2_ntaps = cfg->fir2_ntaps;
        int rc;
        if ((rc = dalloc(&e->st.agc_gain, C)) || (rc = dalloc(&e->st.agc_power, C)) ||
            (rc = dalloc(&e->st.lf_gain, C)) || (rc = dalloc(&e->st.lf_power, C)) ||
            (rc = dalloc(&e->st.sql_state, C)) || (rc = dalloc(&e->st.sql_prev, C)) ||
            (rc = dalloc(&e->d_fm[0], C)) || (rc = dalloc(&e->d_fm[1], C)) || (rc = dalloc(&e->d_mod, C)))
            return rc;
        {
            std::vector<uint8_t> mod(C, SDRX_MOD_AM);
            for (int c = 0; c < C && cfg->modulation; ++c) {
                if (cfg->modulation[c % Cs] > SDRX_MOD_FM) return fail(SDRX_ERR_ARG, "modulation must be SDRX_MOD_AM or SDRX_MOD_FM");
                mod[c] = cfg->modulation[c % Cs];
            }
            if ((rc = upload(e->d_mod, mod.data(), C))) return rc;
            e->st.mod = e->d_mod;
        }
        std::vector<float> win(kChunk);
        sdrx_make_window(win.data());
        std::vector<float2> tw(kChunk / 2);
        for (int k = 0; k < kChunk / 2; ++k) {      // oracle/fft512.c: double, rounded once
            const double a = -2.0 * M_PI * (double)k / (double)kChunk;
            tw[k] = make_float2((float)cos(a), (float)sin(a));
        }
        std::vector<uint16_t> rev(kChunk);
        for (int i = 0; i < kChunk; ++i) {
            unsigned r = 0;
            for (int b = 0; b < 9; ++b) if (i & (1 << b)) r |= 1u << (8 - b);
            rev[i] = (uint16_t)r;
        }
        std::vector<float> sql(C, 9.0f);
        std::vector<int> pos(C, 0);
        for (int c = 0; c < C; ++c) {
            if (cfg->sql_level) sql[c] = cfg->sql_level[c % Cs];
            if (cfg->pos) pos[c] = cfg->pos[c % Cs];
        }
        std::vector<float> c_adj(cfg->fir2_taps, cfg->fir2_taps + cfg->fir2_ntaps);
        if (cfg->lf_gain_db != 0.0f)            // FIR2::setGain, fir.hpp:160-170 (the ctor leaves c_adj_ = c_)
            for (auto &c : c_adj) c = c * std::pow(10.0f, cfg->lf_gain_db / 20.0f);
        if ((rc = dalloc(&e->d_window, kChunk)) || (rc = upload(e->d_window, win.data(), kChunk)) ||
            (rc = dalloc(&e->d_twiddle, kChunk / 2)) || (rc = upload(e->d_twiddle, tw.data(), kChunk / 2)) ||
            (rc = dalloc(&e->d_bitrev, kChunk)) || (rc = upload(e->d_bitrev, rev.data(), kChunk)) ||
            (rc = dalloc(&e->d_shape, kChunk)) || (rc = upload(e->d_shape, cfg->passband_shape, kChunk)) ||
            (rc = dalloc(&e->d_ramp_up, kChunk)) || (rc = upload(e->d_ramp_up, cfg->ramp_up, kChunk)) ||
            (rc = dalloc(&e->d_ramp_down, kChunk)) || (rc = upload(e->d_ramp_down, cfg->ramp_down, kChunk)) ||
            (rc = dalloc(&e->d_sql_level, C)) || (rc = upload(e->d_sql_level, sql.data(), C)) ||
            (rc = dalloc(&e->d_pos, C)) || (rc = upload(e->d_pos, pos.data(), C)) ||
            (rc = dalloc(&e->d_c_adj, c_adj.size())) || (rc = upload(e->d_c_adj, c_adj.data(), c_adj.size())) ||
            (rc = dalloc(&e->d_fir_hist, (size_t)NS * 2 * (size_t)std::max(cfg->fir2_ntaps - 1, 1))))
            return rc;
        if ((rc = dalloc(&e->d_ticket, 2 * (size_t)NS))) return rc;
        {
            // Where boundary A is short (table GEMM / polyphase) the recurrence CTAs ask for (nearly) all of an SM's shared memory and
            // run deep tile rings: nothing else -- detector, envelope, mix, fold CTAs -- is placed beside a dependent chain whose issue
            // slots it would share.  Under the time-domain bank they keep the three-tile rings that fit the SMs its schedule leaves.
            const char *ex = getenv("SDRX_RECUR_EXCL");
            const bool deep = ((e->use_tgemm || e->poly) && (!ex || atoi(ex) != 0)) || (ex && atoi(ex) == 2);   // 2: in every mode (experiment)
            e->agc_bufs = deep ? kAgcBufs : 3;
            e->lf_bufs = deep ? kLfBufs : 3;
            e->agc_smem = deep ? kRecurExclusiveSmem : agc_smem_bytes(3);
            e->lf_smem = deep ? kRecurExclusiveSmem : std::max<size_t>(lf_smem_bytes(3), 120 * 1024);
        }
        if ((rc = raise_dyn_smem((const void *)k_chan_agc, e->agc_smem)) || (rc = raise_dyn_smem((const void *)k_chan_lfagc, e->lf_smem)))
            return rc;
        if ((rc = dalloc(&e->d_agc_iq[0], (size_t)e->Cpad * kChunk)) || (rc = dalloc(&e->d_agc_iq[1], (size_t)e->Cpad * kChunk)) || (rc = dalloc(&e->d_env[0], (size_t)e->Cpad * kChunk)) ||
            (rc = dalloc(&e->d_env[1], (size_t)e->Cpad * kChunk)))
            return rc;
        e->tb = DemodTables{e->d_window, e->d_twiddle, e->d_bitrev, e->d_shape, e->d_ramp_up, e->d_ramp_down};
        if ((size_t)C * 4 > 160 * 1024) return fail(SDRX_ERR_UNSUPPORTED, "more than 40960 channels per engine");
        { int rc2 = raise_dyn_smem((const void *)k_mix_fir2_pcm, mix_smem_bytes(e->Cpad, cfg->fir2_ntaps)); if (rc2) return rc2; }
        if (S >= 1) {       // the per-channel tail kernel stages the last stage's input row in shared memory when fused
            size_t row = (size_t)(e->H[S - 1] - 1) + (size_t)kChunk * e->m[S - 1];
            e->tail_fusable = row * sizeof(float2) <= 100 * 1024 && (e->H[S - 1] - 1) / 2 + 1 <= kTailMaxHalf;
            if (e->tail_fusable)
                { int rc2 = raise_dyn_smem((const void *)k_chan_head<kHeadFir>, sizeof(float2) * row); if (rc2) return rc2; }
            e->tail_fast = e->tail_fusable && e->H[S - 1] == 67 && e->m[S - 1] == 5 && e->chiq_stride % 2 == 0;
            if (e->tail_fast) {
                const float *h = e->taps.data() + e->tap_off[S - 1];
                for (int i = 0; i < 67; ++i) e->tail_taps.h[i] = make_float2(h[i], h[i]);
                { int rc2 = raise_dyn_smem((const void *)k_chan_fir_fast<67, 5>, fir_fast_smem(67, 5, kChunk * 5)); if (rc2) return rc2; }
                CU(cudaFuncSetAttribute(k_chan_fir_fast<67, 5>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                        (int)cudaSharedmemCarveoutMaxShared));
                if (e->use_fast) {
                    int rc2 = dalloc(&e->d_stage_alt, (size_t)e->stage_stride[S - 1] * C);
                    if (rc2) return rc2;
                }
            }
        }
    }
    // Every kernel of the block asks for the same (maximum) shared-memory carve-out: an SM only co-schedules CTAs of
    // different kernels when they agree on the L1/shared split, and the side-stream kernels are meant to run next to
    // the channel bank, which needs all of it.
    {
        const int mx = (int)cudaSharedmemCarveoutMaxShared;
        CU(cudaFuncSetAttribute(k_chan_agc, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_chan_env, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_chan_lfagc, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_mix_fir2_pcm, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_chan_head<kHeadFir>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_chan_head<kHeadDetect>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_prep<SDRX_FMT_U8>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_prep<SDRX_FMT_S16>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
        CU(cudaFuncSetAttribute(k_prep<SDRX_FMT_F32>, cudaFuncAttributePreferredSharedMemoryCarveout, mx));
    }
    if (NS > 1 && !e->use_fast && !e->use_tgemm)
        return fail(SDRX_ERR_UNSUPPORTED, "a multi-stream engine needs the specialised channel bank (the 10 MS/s chain, translator tables, no --ftfir / FORCE_GENERIC)");
    // dalloc's memsets were queued on the legacy default stream, which the engine's non-blocking streams never wait for
    CU(cudaDeviceSynchronize());
    return reset_state(e);
}

}  // namespace

extern "C" int sdrx_engine_create(const sdrx_engine_config *cfg, sdrx_engine **out) {
    if (!cfg || !out) return fail(SDRX_ERR_ARG, "null argument");
    *out = nullptr;
    if (cfg->n_channels < 1) return fail(SDRX_ERR_ARG, "n_channels < 1");
    if (cfg->n_stages < 1 || cfg->n_stages > SDRX_MAX_STAGES) return fail(SDRX_ERR_ARG, "n_stages out of range");
    if (!cfg->m || !cfg->ntaps || !cfg->taps) return fail(SDRX_ERR_ARG, "stage list missing");
    if (cfg->block_len < 1) return fail(SDRX_ERR_ARG, "block_len < 1");
    if (cfg->n_streams < 0 || cfg->n_streams > 4096) return fail(SDRX_ERR_ARG, "n_streams out of range");
    if (cfg->n_streams > 1 && cfg->n_channels % 16 != 0)
        return fail(SDRX_ERR_UNSUPPORTED, "a multi-stream engine needs a multiple of 16 channels per stream");
    for (int s = 0; s < cfg->n_stages; ++s) {
        if (cfg->m[s] < 1) return fail(SDRX_ERR_ARG, "stage m < 1");
        // msd.hpp:169-170 asserts a non-empty, odd-length (folded FIR) coefficient vector
        if (cfg->ntaps[s] < 1 || (cfg->ntaps[s] & 1) == 0) return fail(SDRX_ERR_ARG, "stage taps must be odd in number");
        if (cfg->ntaps[s] > 1025) return fail(SDRX_ERR_UNSUPPORTED, "more than 1025 taps in a stage");
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        cudaGetLastError();
        return fail(SDRX_ERR_CUDA, "no CUDA device: sdrx_b200 has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= ndev) return fail(SDRX_ERR_ARG, "device ordinal out of range");
    DevGuard g(cfg->device);
    if (!g.ok) return fail(SDRX_ERR_CUDA, "cudaSetDevice failed");
    sdrx_engine *e = new sdrx_engine();
    int rc = create_impl(cfg, e);
    if (rc != SDRX_OK) {
        std::string keep = last_error_ref();
        sdrx_engine_destroy(e);
        last_error_ref() = keep;
        return rc;
    }
    *out = e;
    return SDRX_OK;
}

extern "C" int sdrx_engine_reset(sdrx_engine *e) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    DevGuard g(e->dev);
    CU(cudaStreamSynchronize(e->s_copy));
    CU(cudaStreamSynchronize(e->s_prep));
    return reset_state(e);
}

// ------------------------------------------------------------------------------------------------
// One block through the chain.
namespace {

enum SrcKind { SRC_HOST, SRC_DEVICE, SRC_NONE };

// debug timeline: kinds in the order a block walks them
enum TlKind { TL_START = 0, TL_PREP, TL_BANK, TL_TAIL, TL_DET, TL_AGC, TL_ENV, TL_LFAGC, TL_MIX, TL_N };
inline void tl_mark(sdrx_engine *e, int kind, cudaStream_t st) {
    if (!e->tl_on) return;
    cudaEventRecord(e->tl_ev[e->blocks % sdrx_engine::kTlBlocks][kind], st);
}

// Stages [s_begin, s_end) of the chain.  `hist` = buffer index of call-relative input 0, `pos0` its
// trans_pos_, `valid_lo` the first buffer index that holds stream samples.
template <int FMT>
int launch_bank_generic(sdrx_engine *e, const uint8_t *raw, int hist, int pos0, int valid_lo, const int *first,
                        const int *n_out, OutSet &o, int s_begin, int s_end, bool carry_last = false) {
    const int S = e->S;
    for (int s = s_begin; s < s_end; ++s) {
        float2 *dst;
        long dst_stride;
        int dst_off;
        if (s + 1 < S) { dst = e->d_stage_in[s + 1]; dst_stride = e->stage_stride[s + 1]; dst_off = e->H[s + 1] - 1; }
        else           { dst = o.ch_iq;              dst_stride = e->chiq_stride;         dst_off = 0; }
        if (n_out[s] > 0) {
            if (s == 0) {
                dim3 grid(div_up(n_out[0], 8), e->Cpad / 32), block(32, 8);
                if (e->ftfir)
                    k_ftfir_generic<FMT><<<grid, block, 0, e->s_main>>>(
                        raw, hist, e->d_hk, e->ft_sets, e->ft_k, e->Cpad, e->C, e->d_has_tr, e->d_taps + e->tap_off[0],
                        e->H[0], e->m[0], first[0], n_out[0], dst, dst_stride, dst_off, valid_lo);
                else
                k_mix_fir_generic<FMT><<<grid, block, 0, e->s_main>>>(
                    raw, hist, e->d_table_t, e->Cpad, e->C, e->N, pos0, e->d_has_tr,
                    e->d_taps + e->tap_off[0], e->H[0], e->m[0], first[0], n_out[0], dst, dst_stride, dst_off,
                    valid_lo);
            } else {
                dim3 grid(div_up(n_out[s], 128), e->C);
                k_fir_generic<<<grid, 128, 0, e->s_main>>>(e->d_stage_in[s], e->stage_stride[s], e->H[s] - 1,
                                                           e->d_taps + e->tap_off[s], e->H[s], e->m[s], first[s],
                                                           n_out[s], dst, dst_stride, dst_off);
            }
            ++e->launches;
        }
        // slide the delay line of stage s (s >= 1; stage 0's is the raw ping-pong prefix)
        if (s >= 1 && e->H[s] > 1) {
            const int n_in = n_out[s - 1];
            if (n_in > 0) {
                k_carry_hist<<<e->C, 256, 0, e->s_main>>>(e->d_stage_in[s], e->stage_stride[s], e->H[s] - 1, n_in);
                ++e->launches;
            }
        }
    }
    // the delay line of the stage after the last one run (materialisation of stages 1-2 only)
    if (carry_last && s_end < S && e->H[s_end] > 1 && n_out[s_end - 1] > 0) {
        k_carry_hist<<<e->C, 256, 0, e->s_main>>>(e->d_stage_in[s_end], e->stage_stride[s_end], e->H[s_end] - 1, n_out[s_end - 1]);
        ++e->launches;
    }
    CU(cudaGetLastError());
    return SDRX_OK;
}

int launch_bank_generic_fmt(sdrx_engine *e, int fmt, const uint8_t *raw, int hist, int pos0, int valid_lo, const int *first,
                            const int *n_out, OutSet &o, int s_begin, int s_end, bool carry_last = false) {
    switch (fmt) {
        case SDRX_FMT_U8:  return launch_bank_generic<SDRX_FMT_U8>(e, raw, hist, pos0, valid_lo, first, n_out, o, s_begin, s_end, carry_last);
        case SDRX_FMT_S16: return launch_bank_generic<SDRX_FMT_S16>(e, raw, hist, pos0, valid_lo, first, n_out, o, s_begin, s_end, carry_last);
        default:           return launch_bank_generic<SDRX_FMT_F32>(e, raw, hist, pos0, valid_lo, first, n_out, o, s_begin, s_end, carry_last);
    }
}

// The specialised bank keeps no stage 1-2 delay lines (it recomputes from the raw history).  Before the
// generic kernels take over (ragged block) or the state is saved, rebuild them by running generic
// stages 0-1 over the last 625 samples in front of buffer index `end_idx`: the last 18 / 20 outputs of
// stages 1 / 2 depend on nothing older.  Stage phases are zero whenever this is needed.
// (The table GEMM keeps no delay line at all: there the replay covers H_last + 7 inputs of the last stage and runs every
// stage but the last, e->mat_len / e->mat_stages.)
int materialize_generic(sdrx_engine *e, int fmt, const uint8_t *buf, long end_idx, uint64_t total_before_end) {
    if (!e->generic_stale) return SDRX_OK;
    const int kMaterialize = e->mat_len;
    const int hist = (int)(end_idx - kMaterialize);
    const int valid_lo = (int)std::max<long>(0, end_idx - (long)std::min<uint64_t>(total_before_end, (uint64_t)end_idx));
    long pos = ((long)e->trans_pos - kMaterialize) % e->N;
    if (pos < 0) pos += e->N;
    int first[SDRX_MAX_STAGES] = {0}, n_out[SDRX_MAX_STAGES] = {0};
    long n_in = kMaterialize;
    for (int s = 0; s < e->mat_stages; ++s) { first[s] = e->m[s] - 1; n_out[s] = (int)(n_in / e->m[s]); n_in = n_out[s]; }
    OutSet dummy{};
    int rc = launch_bank_generic_fmt(e, fmt, buf, hist, (int)pos, valid_lo, first, n_out, dummy, 0, e->mat_stages, true);
    if (rc) return rc;
    e->generic_stale = false;
    return SDRX_OK;
}

template <int FMT>
int launch_bank_fast(sdrx_engine *e, const uint8_t *raw, int valid_lo, unsigned n, OutSet &o, const FastSchedule **deferred, int tail_half) {
    float2 *dst; long stride; int off;
    if (e->S > 3) { dst = e->d_stage_in[3]; stride = e->stage_stride[3]; off = e->H[3] - 1; }
    else          { dst = o.ch_iq;          stride = e->chiq_stride;     off = 0; }
    return fast_launch<FMT>(e->fast, e->s_main, raw, e->raw_hist, valid_lo, e->N, (int)e->trans_pos, n,
                            dst, stride, off, &e->launches,
                            e->timing ? e->ev_b0 : nullptr, e->timing ? e->ev_b1 : nullptr, deferred, tail_half, (long)e->raw_stream_bytes);
}
int launch_bank_fast_fmt(sdrx_engine *e, int fmt, const uint8_t *raw, int valid_lo, unsigned n, OutSet &o,
                         const FastSchedule **deferred = nullptr, int tail_half = 0) {
    switch (fmt) {
        case SDRX_FMT_U8:  return launch_bank_fast<SDRX_FMT_U8>(e, raw, valid_lo, n, o, deferred, tail_half);
        case SDRX_FMT_S16: return launch_bank_fast<SDRX_FMT_S16>(e, raw, valid_lo, n, o, deferred, tail_half);
        default:           return launch_bank_fast<SDRX_FMT_F32>(e, raw, valid_lo, n, o, deferred, tail_half);
    }
}

// power partials of the new block (when boundary B will run) + the raw delay line, one launch
int launch_prep(sdrx_engine *e, cudaStream_t st, int fmt, uint8_t *buf, const uint8_t *prev, unsigned n, bool power, int hist_bytes, long prev_bytes, float *partial) {
    const int np = power ? kPwrPartials : 0;
    const dim3 grid(np + div_up(hist_bytes, 256), e->NS);
    if (grid.x == 0) return SDRX_OK;
    const long sb = (long)e->raw_stream_bytes;
    switch (fmt) {
        case SDRX_FMT_U8:  k_prep<SDRX_FMT_U8><<<grid, 256, 0, st>>>(buf, e->raw_hist, (int)n, partial, np, buf, prev, prev_bytes, hist_bytes, sb); break;
        case SDRX_FMT_S16: k_prep<SDRX_FMT_S16><<<grid, 256, 0, st>>>(buf, e->raw_hist, (int)n, partial, np, buf, prev, prev_bytes, hist_bytes, sb); break;
        default:           k_prep<SDRX_FMT_F32><<<grid, 256, 0, st>>>(buf, e->raw_hist, (int)n, partial, np, buf, prev, prev_bytes, hist_bytes, sb); break;
    }
    ++e->launches;
    CU(cudaGetLastError());
    return SDRX_OK;
}

// Boundary B on the chunk in o.ch_iq (output set `idx`).  Only the last MSD stage (when `fused`) runs on the main
// stream; the squelch detector, the recurrences and the mix follow on the side stream, where their CTAs sit next
// to the next block's channel-bank CTAs (none of them needs more than the 9 KB of shared memory the bank leaves).
int launch_mix(sdrx_engine *e, cudaStream_t st, OutSet &o, const float *pwr_partial, int n_partial, unsigned n_samples,
               float *pwr, const float *ext_parts, int n_ext, int do_fir) {
    // sdrx_engine_finish_mix (ext_parts) may run next to a block's own mix kernel: it takes its own arrival counter
    k_mix_fir2_pcm<<<dim3(kMixCtas, ext_parts ? 1 : e->NS), kMixThreads, mix_smem_bytes(e->Cpad, e->fir2_ntaps), st>>>(
        o.audio, e->Cpad, o.mixed, e->d_pos, e->Cs, e->d_c_adj, e->fir2_ntaps, e->d_fir_hist, o.mix_pre, o.mix, o.pcm, pwr_partial,
        n_partial, n_samples, pwr, ext_parts, n_ext, do_fir, o.audio_idx, o.n_audio, e->d_ticket + (ext_parts ? e->NS : 0));
    ++e->launches;
    CU(cudaGetLastError());
    return SDRX_OK;
}

int launch_demod(sdrx_engine *e, int idx, bool have_power, unsigned n_samples, const TailFirArgs *fused = nullptr, int side_par = -1) {
    OutSet &o = e->out[idx];
    const TailSnap snap{o.sql, o.agc_gain, o.lf_gain};
    if (fused && side_par >= 0) {
        // the specialised last stage under the next block's bank: main (bank done) -> detector stream -> recurrences
        CU(cudaEventRecord(e->ev_head_done, e->s_main));
        CU(cudaStreamWaitEvent(e->s_side, e->ev_head_done, 0));
        float2 *rows_next = (fused->stage_in == e->d_stage_alt) ? e->d_stage_in[e->S - 1] : e->d_stage_alt;
        k_chan_fir_fast<67, 5><<<e->C, kHeadThreads, fir_fast_smem(67, 5, fused->n_in), e->s_side>>>(*fused, rows_next, e->tail_taps, o.ch_iq,
                                                                                                    e->chiq_stride);
        ++e->launches;
        CU(cudaGetLastError());
        tl_mark(e, TL_TAIL, e->s_side);
        CU(cudaEventRecord(e->ev_tail_done[side_par], e->s_side));
        e->tail_pending[side_par] = true;
        e->last_tail_par = side_par;
        CU(cudaStreamWaitEvent(e->s_recur, e->ev_tail_done[side_par], 0));
    } else {
        if (fused) {
            const size_t dyn = sizeof(float2) * (size_t)(fused->H - 1 + fused->n_in);
            if (e->tail_fast && fused->n_in == kChunk * 5 && fused->first >= 0 && fused->first < 5)
                k_chan_fir_fast<67, 5><<<e->C, kHeadThreads, fir_fast_smem(67, 5, fused->n_in), e->s_main>>>(*fused, fused->stage_in, e->tail_taps,
                                                                                                            o.ch_iq, e->chiq_stride);
            else
                k_chan_head<kHeadFir><<<e->C, kHeadThreads, dyn, e->s_main>>>(*fused, o.ch_iq, e->chiq_stride, e->st, e->tb, e->d_sql_level,
                                                                              o.snr, e->d_gate[idx], snap);
            ++e->launches;
        }
        CU(cudaGetLastError());
        if (!e->use_tc) tl_mark(e, TL_TAIL, e->s_main);
        CU(cudaEventRecord(e->ev_head_done, e->s_main));
        CU(cudaStreamWaitEvent(e->s_side, e->ev_head_done, 0));
        CU(cudaStreamWaitEvent(e->s_recur, e->ev_head_done, 0));
    }
    k_chan_head<kHeadDetect><<<e->C, kHeadThreads, 0, e->s_side>>>(TailFirArgs{}, o.ch_iq, e->chiq_stride, e->st, e->tb,
                                                                   e->d_sql_level, o.snr, e->d_gate[idx], snap);
    tl_mark(e, TL_DET, e->s_side);
    CU(cudaEventRecord(e->ev_det_done[idx], e->s_side));
    // the AGC runs whatever the squelch state (sdrx.cpp:457): next to the detector, not behind it -- and on a stream of its
    // own, so that block b + 1's AGC recurrence runs next to block b's LfAGC recurrence (gain-adjusted IQ in a ping-pong)
    const unsigned par = e->agc_par++ & 1u;
    // ... where boundary A is short (table GEMM / polyphase) and boundary B sets the step.  Under the time-domain bank the
    // recurrences have the three SMs its schedule leaves free and a whole bank's time: one stream, so that two blocks'
    // recurrence CTAs never hold more SMs than that when the next bank launch needs its own
    cudaStream_t s_lf = (e->use_tgemm || e->poly) ? e->s_lf : e->s_recur;
    k_chan_agc<<<recur_ctas(e->C), kRecurThreads, e->agc_smem, e->s_recur>>>(o.iq_map, e->st, e->agc, e->d_agc_iq[par], e->Cpad, snap, e->C, e->agc_bufs);
    tl_mark(e, TL_AGC, e->s_recur);
    // the envelope follows the AGC on its stream (it needs the detector's gate); the LfAGC recurrence has a stream to itself
    if (e->env_pending[par]) CU(cudaStreamWaitEvent(e->s_recur, e->ev_env_done[par], 0));   // the LfAGC of two blocks ago read d_env[par]
    CU(cudaStreamWaitEvent(e->s_recur, e->ev_det_done[idx], 0));
    ChannelState st_env = e->st;
    st_env.fm_prev = e->d_fm[e->fm_cur];
    k_chan_env<<<dim3(e->Cpad / 32, kChunk / 4), dim3(32, 4), 0, e->s_recur>>>(e->d_agc_iq[par], e->Cpad, st_env, e->d_gate[idx], e->d_env[par],
                                                                                e->d_fm[e->fm_cur ^ 1], e->C);
    e->fm_cur ^= 1;
    tl_mark(e, TL_ENV, e->s_recur);
    CU(cudaEventRecord(e->ev_agc_done[par], e->s_recur));
    CU(cudaStreamWaitEvent(s_lf, e->ev_agc_done[par], 0));
    k_chan_lfagc<<<recur_ctas(e->C), kRecurThreads, e->lf_smem, s_lf>>>(e->d_env[par], e->Cpad, e->st, e->d_gate[idx], e->tb, e->lf, e->lf_active, o.audio, snap,
                                                        e->C, o.mixed, e->lf_bufs);
    tl_mark(e, TL_LFAGC, s_lf);
    CU(cudaEventRecord(e->ev_env_done[par], s_lf));          // d_env[par] (and with it d_agc_iq[par]) are free again
    e->env_pending[par] = true;
    CU(cudaEventRecord(e->ev_recur_done[idx], s_lf));
    CU(cudaStreamWaitEvent(e->s_mix, e->ev_recur_done[idx], 0));
    e->launches += 4;
    int rc = launch_mix(e, e->s_mix, o, e->d_pwr_partial[idx], have_power ? kPwrPartials : 0, n_samples,
                        have_power ? o.pwr : nullptr, nullptr, 0, (e->flags & SDRX_FLAG_PARTIAL_MIX) ? 0 : 1);
    if (rc) return rc;
    if (e->timing) CU(cudaEventRecord(e->ev_t2, e->s_mix));
    tl_mark(e, TL_MIX, e->s_mix);
    CU(cudaEventRecord(e->ev_side_done[idx], e->s_mix));
    e->side_pending[idx] = true;
    return SDRX_OK;
}

// the main stream may not touch output set `idx` (or its power partials / gate) before the side stream is done with it
int wait_side(sdrx_engine *e, int idx) {
    if (e->side_pending[idx]) CU(cudaStreamWaitEvent(e->s_main, e->ev_side_done[idx], 0));
    if (e->fetch_pending[idx]) CU(cudaStreamWaitEvent(e->s_main, e->ev_fetch_done[idx], 0));
    return SDRX_OK;
}

// device-visible alias of a pinned host pointer (nullptr: pageable / unknown)
void *device_view(const void *host) {
    if (!host) return nullptr;
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (attr.type != cudaMemoryTypeHost && attr.type != cudaMemoryTypeManaged) return nullptr;
    return attr.devicePointer;
}

// `blocking`: the caller synchronises `st` itself afterwards and pageable sparse buffers are allowed (staged through a
// device scratch array and copied after the count is known); otherwise the sparse fields must be pinned.
int copy_outputs(sdrx_engine *e, const OutSet &o, const sdrx_block_outputs *out, cudaStream_t st, bool blocking) {
    const size_t C = (size_t)e->C;
    auto d2h = [&](void *dst, const void *src, size_t bytes) -> cudaError_t {
        return dst ? cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    const dim3 xgrid((unsigned)std::min<int>(e->Cs, 128), (unsigned)e->NS);
    const size_t NS = (size_t)e->NS;
    float *&rows_scratch = e->d_rows_scratch[blocking ? 0 : 1];
    auto scratch = [&]() -> int {
        if (!rows_scratch) CU(cudaMalloc((void **)&rows_scratch, sizeof(float) * C * kChunk));
        return SDRX_OK;
    };
    if (out->audio) {             // the dense [C][512] array: rows out of the sample-major device array, then one copy
        { int rc = scratch(); if (rc) return rc; }
        k_export_audio<<<xgrid, 128, 0, st>>>(o.audio, e->Cpad, nullptr, nullptr, e->Cs, rows_scratch, nullptr, nullptr, e->Cs);
        ++e->launches;
        CU(cudaGetLastError());
        CU(d2h(out->audio, rows_scratch, C * kChunk * sizeof(float)));
    }
    CU(d2h(out->sql_state, o.sql, C));
    CU(d2h(out->snr, o.snr, C * sizeof(float)));
    CU(d2h(out->mix_pre, o.mix_pre, NS * 2 * kChunk * sizeof(float)));
    CU(d2h(out->mix, o.mix, NS * 2 * kChunk * sizeof(float)));
    CU(d2h(out->pcm, o.pcm, NS * 2 * kChunk * sizeof(int16_t)));
    CU(d2h(out->pwr_dbfs, o.pwr, NS * sizeof(float)));
    CU(d2h(out->agc_gain, o.agc_gain, C * sizeof(float)));
    CU(d2h(out->lf_gain, o.lf_gain, C * sizeof(float)));
    if (out->audio_rows || out->audio_chan || out->n_audio) {
        float *rows = (float *)device_view(out->audio_rows);
        int *chan = (int *)device_view(out->audio_chan);
        int *cnt = (int *)device_view(out->n_audio);
        const bool direct = (!out->audio_rows || rows) && (!out->audio_chan || chan) && (!out->n_audio || cnt);
        if (direct) {
            k_export_audio<<<xgrid, 128, 0, st>>>(o.audio, e->Cpad, o.audio_idx, o.n_audio, 0, rows, chan, cnt, e->Cs);
            ++e->launches;
            CU(cudaGetLastError());
        } else {
            if (!blocking) return fail(SDRX_ERR_ARG, "fetch_outputs: audio_rows / audio_chan / n_audio must be pinned (sdrx_host_alloc)");
            if (NS > 1) return fail(SDRX_ERR_UNSUPPORTED, "multi-stream engine: audio_rows / audio_chan / n_audio must be pinned (sdrx_host_alloc)");
            int n = 0;
            CU(cudaMemcpyAsync(&n, o.n_audio, sizeof(int), cudaMemcpyDeviceToHost, st));
            CU(cudaStreamSynchronize(st));
            if (out->n_audio) *out->n_audio = n;
            if (out->audio_chan && n > 0) CU(d2h(out->audio_chan, o.audio_idx, sizeof(int) * (size_t)n));
            if (out->audio_rows && n > 0) {
                { int rc = scratch(); if (rc) return rc; }
                k_export_audio<<<xgrid, 128, 0, st>>>(o.audio, e->Cpad, o.audio_idx, o.n_audio, 0, rows_scratch, nullptr, nullptr, e->Cs);
                ++e->launches;
                CU(cudaGetLastError());
                CU(d2h(out->audio_rows, rows_scratch, sizeof(float) * (size_t)n * kChunk));
            }
        }
    }
    return SDRX_OK;
}

int enqueue_block(sdrx_engine *e, const void *src, SrcKind kind, int fmt, unsigned n) {
    if (fmt != SDRX_FMT_U8 && fmt != SDRX_FMT_S16 && fmt != SDRX_FMT_F32) return fail(SDRX_ERR_ARG, "unknown IQ format");
    if (n > e->block_len) return fail(SDRX_ERR_ARG, "block longer than block_len");
    if (e->fmt >= 0 && e->fmt != fmt && e->pushes > 0)
        return fail(SDRX_ERR_STATE, "wire format changed mid-stream (the carried raw history is format specific); reset first");
    e->fmt = fmt;
    const int bps = raw_bytes_per_sample(fmt);
    const int cur = (int)(e->pushes % kNBuf), prv = (cur + kNBuf - 1) % kNBuf;   // raw ring: this block's buffer, the previous block's
    const int bpar = (int)(e->pushes & 1);                     // ping-pong of the tensor-core fold's B planes
    const int oc = (int)(e->blocks % e->ring);                 // output set of this block
    uint8_t *buf = e->d_raw[cur];
    const size_t hist_bytes = (size_t)e->raw_hist * bps;


    // ---- stage bookkeeping (MSD::S::isn_, msd.hpp:240-244): which inputs complete an output
    int first[SDRX_MAX_STAGES], n_out[SDRX_MAX_STAGES], phase_in[SDRX_MAX_STAGES] = {0};
    long n_in = n;
    for (int s = 0; s < e->S; ++s) phase_in[s] = e->phase[s];
    for (int s = 0; s < e->S; ++s) {
        first[s] = e->m[s] - 1 - e->phase[s];
        n_out[s] = (int)((e->phase[s] + n_in) / e->m[s]);
        e->phase[s] = (int)((e->phase[s] + n_in) % e->m[s]);
        n_in = n_out[s];
    }
    const unsigned out_len = (unsigned)n_in;
    OutSet &o = e->out[oc];
    const bool demod = !(e->flags & SDRX_FLAG_NO_DEMOD) && out_len == (unsigned)kChunk;
    // ---- ingest: pinned, double buffered, on the copy stream
    // buffer `cur` was last read by block b-2's raw carry (and before that by block b-3's kernels)
    CU(cudaStreamWaitEvent(e->s_copy, e->ev_raw_free[cur], 0));
    // ... and on the prep stream by block b-2's k_prep (the delay-line tail) with block b-3's fold queued in front of it: both a
    // whole block back, so the H2D of block b never waits for anything of block b-1
    { const int b2 = (cur + 1) % kNBuf;                        // the buffer of block b-2, whose k_prep read this buffer's tail
      if (e->prep_pending[b2]) CU(cudaStreamWaitEvent(e->s_copy, e->ev_prep[b2], 0)); }
    if (n > 0) {
        // one stream: a plain copy; several: the caller's [n_streams][n] blocks land behind each stream's history prefix
        const size_t row = (size_t)n * bps;
        auto copy_in = [&](const void *from, cudaMemcpyKind k) -> cudaError_t {
            if (e->NS == 1) return cudaMemcpyAsync(buf + hist_bytes, from, row, k, e->s_copy);
            // one plain copy per stream: a 2-D copy of 64 rows of 2.56 MB failed on this driver ("unspecified launch
            // failure" device to device; 24 rows worked), so the 2-D form is not used for any direction
            for (int q = 0; q < e->NS; ++q) {
                const cudaError_t r = cudaMemcpyAsync(buf + hist_bytes + (size_t)q * e->raw_stream_bytes, (const uint8_t *)from + (size_t)q * row,
                                                      row, k, e->s_copy);
                if (r != cudaSuccess) return r;
            }
            return cudaSuccess;
        };
        if (kind == SRC_DEVICE) {
            CU(copy_in(src, cudaMemcpyDeviceToDevice));
        } else {
            cudaPointerAttributes attr{};
            bool pinned = cudaPointerGetAttributes(&attr, src) == cudaSuccess &&
                          (attr.type == cudaMemoryTypeHost || attr.type == cudaMemoryTypeManaged);
            cudaGetLastError();
            const void *from = src;
            if (!pinned) {
                if (!e->h_stage[cur]) CU(cudaMallocHost((void **)&e->h_stage[cur], (size_t)e->block_len * 8 * e->NS));
                if (e->stage_busy[cur]) CU(cudaEventSynchronize(e->ev_copied[cur]));
                memcpy(e->h_stage[cur], src, row * e->NS);
                e->stage_busy[cur] = true;
                from = e->h_stage[cur];
            }
            CU(copy_in(from, cudaMemcpyHostToDevice));
        }
    }
    // ---- still on the copy stream, off the compute stream's critical path: the raw delay line (last raw_hist
    //      samples of the previous [hist | block] move in front of this block) and, when boundary B will run,
    //      the block power partials.  The partials of set `oc` were last read by block b-kNOut's mix kernel.
    bool tc_block = false;
    const int valid_lo_early = (int)std::max<long>(0, (long)e->raw_hist - (long)std::min<uint64_t>(e->total, (uint64_t)e->raw_hist));
    if (e->use_tgemm && e->use_tc && n > 0) {
        bool pz = true;
        for (int s = 0; s < e->S; ++s) pz = pz && phase_in[s] == 0;
        tc_block = pz && n % (unsigned)e->tg.D == 0 && tgemm_eligible(e->N, (int)out_len) && (int)out_len <= e->tg.Mpad;
    }
    {
        // the copy stream carries only the copies: k_prep and the fold follow on the prep stream, so that the next block's H2D
        // (50 us of PCIe time per 10 MS/s f32 block) runs beside them instead of behind them
        CU(cudaEventRecord(e->ev_h2d[cur], e->s_copy));
        CU(cudaStreamWaitEvent(e->s_prep, e->ev_h2d[cur], 0));
        if (e->side_pending[oc]) CU(cudaStreamWaitEvent(e->s_prep, e->ev_side_done[oc], 0));
        int rc = launch_prep(e, e->s_prep, fmt, buf, e->d_raw[prv], n, demod && n > 0, (int)hist_bytes, (long)e->prev_n * bps,
                             e->d_pwr_partial[oc]);
        if (rc) return rc;
        tl_mark(e, TL_PREP, e->s_prep);
        CU(cudaEventRecord(e->ev_prep[cur], e->s_prep));
        e->prep_pending[cur] = true;
        if (tc_block) {
            // tensor-core table GEMM: the fold of this block runs here, on the copy stream, under the previous block's GEMM
            const long sb = (long)e->raw_stream_bytes;
            switch (fmt) {
                case SDRX_FMT_U8:  rc = tgemm_tc_fold<SDRX_FMT_U8>(e->tg, e->tc, e->s_prep, bpar, buf, sb, e->raw_hist, valid_lo_early, (int)e->trans_pos, (int)out_len, &e->launches); break;
                case SDRX_FMT_S16: rc = tgemm_tc_fold<SDRX_FMT_S16>(e->tg, e->tc, e->s_prep, bpar, buf, sb, e->raw_hist, valid_lo_early, (int)e->trans_pos, (int)out_len, &e->launches); break;
                default:           rc = tgemm_tc_fold<SDRX_FMT_F32>(e->tg, e->tc, e->s_prep, bpar, buf, sb, e->raw_hist, valid_lo_early, (int)e->trans_pos, (int)out_len, &e->launches); break;
            }
            if (rc) return rc;
            tl_mark(e, TL_TAIL, e->s_prep);                      // (timeline: the `tail` column is the fold in this mode)
        }
        CU(cudaEventRecord(e->ev_copied[cur], e->s_prep));
        // every main-stream kernel that read buffer `prv` (the previous block's channelizer) is queued before this point -- recorded
        // BEFORE the main stream is made to wait for this block's prep / fold: behind that wait the event would also stand for them,
        // and the next block's H2D (which waits for it) would run after this block's fold instead of beside it
        CU(cudaEventRecord(e->ev_raw_free[prv], e->s_main));
        CU(cudaStreamWaitEvent(e->s_main, e->ev_copied[cur], 0));
    }
    { int rc = wait_side(e, oc); if (rc) return rc; }       // block b-kNOut's recurrences / mix / read-back used this output set
    if (e->timing) CU(cudaEventRecord(e->ev_t0, e->s_main));
    tl_mark(e, TL_START, e->s_main);

    const int valid_lo = (int)std::max<long>(0, (long)e->raw_hist - (long)std::min<uint64_t>(e->total, (uint64_t)e->raw_hist));
    const bool fast = e->use_fast && n > 0 && n % kUnit == 0 && phase_in[0] == 0 && phase_in[1] == 0 && phase_in[2] == 0;
    if (e->NS > 1 && !e->use_tgemm && !(fast && demod))
        return fail(SDRX_ERR_UNSUPPORTED, "multi-stream engine: blocks must be whole (block_len samples per stream)");
    int rc;
    const bool fuse_tail = fast && demod && e->tail_fusable && e->S == e->fast.stages_fused + 1;
    const int s_last = e->S - 1;
    const bool side_tail = fuse_tail && e->tail_fast && e->d_stage_alt && n_out[s_last - 1] == kChunk * 5 && first[s_last] >= 0 &&
                           first[s_last] < 5;
    int side_par = -1;
    TailFirArgs tail{};
    bool phases_zero = true;
    for (int s = 0; s < e->S; ++s) phases_zero = phases_zero && phase_in[s] == 0;
    const bool tgemm = e->use_tgemm && n > 0 && phases_zero && n % (unsigned)e->tg.D == 0 && tgemm_eligible(e->N, (int)out_len) &&
                       (int)out_len <= e->tg.Mpad;
    if (e->NS > 1 && e->use_tgemm && !tgemm)
        return fail(SDRX_ERR_UNSUPPORTED, "multi-stream engine: blocks must be whole (block_len samples per stream)");
    if (tgemm) {
        cudaEvent_t b0 = e->timing ? e->ev_b0 : nullptr, b1 = e->timing ? e->ev_b1 : nullptr;
        const long sb = (long)e->raw_stream_bytes;
        if (e->use_tc) rc = tgemm_tc_launch(e->tg, e->tc, e->s_main, bpar, (int)out_len, o.ch_iq, e->chiq_stride, &e->launches, b0, b1);
        else switch (fmt) {
            case SDRX_FMT_U8:  rc = tgemm_launch<SDRX_FMT_U8>(e->tg, e->s_main, buf, sb, e->raw_hist, valid_lo, (int)e->trans_pos, (int)out_len, o.ch_iq, e->chiq_stride, &e->launches, b0, b1); break;
            case SDRX_FMT_S16: rc = tgemm_launch<SDRX_FMT_S16>(e->tg, e->s_main, buf, sb, e->raw_hist, valid_lo, (int)e->trans_pos, (int)out_len, o.ch_iq, e->chiq_stride, &e->launches, b0, b1); break;
            default:           rc = tgemm_launch<SDRX_FMT_F32>(e->tg, e->s_main, buf, sb, e->raw_hist, valid_lo, (int)e->trans_pos, (int)out_len, o.ch_iq, e->chiq_stride, &e->launches, b0, b1); break;
        }
        e->generic_stale = true;
        e->bank_timed = e->timing;
    } else if (e->poly) {
        if (!(demod && n % (unsigned)e->poly_D == 0 && phase_in[0] == 0))
            return fail(SDRX_ERR_UNSUPPORTED, "polyphase engine: blocks must be whole (512 outputs per channel)");
        e->bank_timed = e->timing;
        if (e->timing) CU(cudaEventRecord(e->ev_b0, e->s_main));
        switch (fmt) {
            case SDRX_FMT_U8:
                k_poly_channelizer<SDRX_FMT_U8><<<out_len, kPolyThreads, poly_smem_bytes(), e->s_main>>>(
                    buf, e->raw_hist, valid_lo, (int)e->trans_pos, e->poly_D, e->d_hc, e->poly_Lh, e->d_wtab, e->d_bin, e->C, o.ch_iq, e->chiq_stride);
                break;
            case SDRX_FMT_S16:
                k_poly_channelizer<SDRX_FMT_S16><<<out_len, kPolyThreads, poly_smem_bytes(), e->s_main>>>(
                    buf, e->raw_hist, valid_lo, (int)e->trans_pos, e->poly_D, e->d_hc, e->poly_Lh, e->d_wtab, e->d_bin, e->C, o.ch_iq, e->chiq_stride);
                break;
            default:
                k_poly_channelizer<SDRX_FMT_F32><<<out_len, kPolyThreads, poly_smem_bytes(), e->s_main>>>(
                    buf, e->raw_hist, valid_lo, (int)e->trans_pos, e->poly_D, e->d_hc, e->poly_Lh, e->d_wtab, e->d_bin, e->C, o.ch_iq, e->chiq_stride);
                break;
        }
        ++e->launches;
        if (e->timing) CU(cudaEventRecord(e->ev_b1, e->s_main));
        CU(cudaGetLastError());
        rc = SDRX_OK;
    } else if (fast) {
        const FastSchedule *sc = nullptr;
        int tail_half = 0;
        if (side_tail) {
            // row / tails buffer `par` was last read by the last-stage kernel two fused blocks ago
            side_par = e->tail_par;
            tail_half = side_par;
            if (e->tail_pending[side_par]) CU(cudaStreamWaitEvent(e->s_main, e->ev_tail_done[side_par], 0));
        } else if (e->last_tail_par >= 0) {
            // this block works in place on the row buffer whose delay line the last side-stream last stage writes
            CU(cudaStreamWaitEvent(e->s_main, e->ev_tail_done[e->last_tail_par], 0));
        }
        rc = launch_bank_fast_fmt(e, fmt, buf, valid_lo, n, o, fuse_tail ? &sc : nullptr, tail_half);
        if (!rc && fuse_tail) {
            const int s = s_last;
            tail = TailFirArgs{e->d_stage_in[s], e->stage_stride[s], n_out[s - 1], e->H[s], e->m[s], first[s],
                               e->d_taps + e->tap_off[s], sc->n_fixes > 0 ? e->fast.tails_half(tail_half) : nullptr, e->fast.tails_slots,
                               sc->d_fixes, sc->d_fix_range, e->fast.d_chan_group};
            if (side_tail) {
                // from here on "the" last-stage input buffer is the partner: its delay line is what this block's last-stage
                // kernel writes, its new part what the next block's bank writes
                std::swap(e->d_stage_in[s], e->d_stage_alt);
                e->tail_par ^= 1;
            }
        } else if (!rc) {
            rc = launch_bank_generic_fmt(e, fmt, buf, e->raw_hist, (int)e->trans_pos, valid_lo, first, n_out, o, 3, e->S);
        }
        e->generic_stale = true;
        e->bank_timed = e->timing;
    } else {
        e->bank_timed = false;
        if (e->last_tail_par >= 0) CU(cudaStreamWaitEvent(e->s_main, e->ev_tail_done[e->last_tail_par], 0));
        rc = materialize_generic(e, fmt, buf, e->raw_hist, e->total);
        if (!rc) rc = launch_bank_generic_fmt(e, fmt, buf, e->raw_hist, (int)e->trans_pos, valid_lo, first, n_out, o, 0, e->S);
    }
    if (rc) return rc;
    tl_mark(e, TL_BANK, e->s_main);
    if (e->timing) CU(cudaEventRecord(e->ev_t1, e->s_main));

    if (demod) { rc = launch_demod(e, oc, n > 0, n, fuse_tail ? &tail : nullptr, side_par); if (rc) return rc; }
    else if (e->timing) CU(cudaEventRecord(e->ev_t2, e->s_main));
    if (e->timing) e->timing_valid = true;

    e->trans_pos = (e->trans_pos + n) % e->N;
    if (e->ftfir) e->ft_k = (e->ft_k + n_out[0]) % e->ft_sets;
    e->total += n;
    e->prev_n = n;
    e->last_out_len = out_len;
    e->last_demod = demod;
    ++e->blocks;
    ++e->pushes;
    return SDRX_OK;
}

}  // namespace

extern "C" int sdrx_engine_push_block(sdrx_engine *e, const void *raw, int fmt, unsigned n_samples) {
    if (!e || (!raw && n_samples)) return fail(SDRX_ERR_ARG, "null argument");
    DevGuard g(e->dev);
    return enqueue_block(e, raw, SRC_HOST, fmt, n_samples);
}

extern "C" int sdrx_engine_push_block_device(sdrx_engine *e, const void *d_raw, int fmt, unsigned n_samples) {
    if (!e || (!d_raw && n_samples)) return fail(SDRX_ERR_ARG, "null argument");
    DevGuard g(e->dev);
    return enqueue_block(e, d_raw, SRC_DEVICE, fmt, n_samples);
}

extern "C" int sdrx_engine_wait_event(sdrx_engine *e, void *cuda_event) {
    if (!e || !cuda_event) return fail(SDRX_ERR_ARG, "null argument");
    DevGuard g(e->dev);
    CU(cudaStreamWaitEvent(e->s_copy, (cudaEvent_t)cuda_event, 0));
    return SDRX_OK;
}

// The event that completes when the last pushed block's ingest copy (and raw carry) has run: a device-side producer
// may overwrite the buffer it handed to push_block_device after this.
extern "C" void *sdrx_engine_ingest_event(sdrx_engine *e) {
    if (!e || e->pushes == 0) return nullptr;
    return (void *)e->ev_copied[(e->pushes + kNBuf - 1) % kNBuf];
}

extern "C" int sdrx_engine_last_out_len(sdrx_engine *e, unsigned *out_len) {
    if (!e || !out_len) return fail(SDRX_ERR_ARG, "null argument");
    *out_len = e->last_out_len;
    return SDRX_OK;
}

extern "C" int sdrx_engine_sync(sdrx_engine *e) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    DevGuard g(e->dev);
    CU(cudaStreamSynchronize(e->s_copy));
    CU(cudaStreamSynchronize(e->s_prep));
    CU(cudaStreamSynchronize(e->s_main));
    CU(cudaStreamSynchronize(e->s_side));
    CU(cudaStreamSynchronize(e->s_recur));
    CU(cudaStreamSynchronize(e->s_lf));
    CU(cudaStreamSynchronize(e->s_mix));
    CU(cudaStreamSynchronize(e->s_out));
    return SDRX_OK;
}

// Orders the main stream behind everything queued so far on the side stream, so that an event recorded on
// sdrx_engine_stream() afterwards covers the whole of every pushed block (bench timing).
extern "C" int sdrx_engine_join(sdrx_engine *e) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    DevGuard g(e->dev);
    for (int b = 0; b < kNOut; ++b) { int rc = wait_side(e, b); if (rc) return rc; }
    return SDRX_OK;
}

extern "C" int sdrx_engine_get_channel_iq(sdrx_engine *e, float *ch_iq) {
    if (!e || !ch_iq) return fail(SDRX_ERR_ARG, "null argument");
    if (e->blocks == 0) return fail(SDRX_ERR_STATE, "no block pushed yet");
    DevGuard g(e->dev);
    const OutSet &o = e->out[(e->blocks - 1) % e->ring];
    { int rc = wait_side(e, (int)((e->blocks - 1) % e->ring)); if (rc) return rc; }   // the last stage may run on a side stream
    if (e->last_out_len > 0)
        CU(cudaMemcpy2DAsync(ch_iq, (size_t)e->last_out_len * sizeof(float2), o.ch_iq, (size_t)e->chiq_stride * sizeof(float2),
                             (size_t)e->last_out_len * sizeof(float2), (size_t)e->C, cudaMemcpyDeviceToHost, e->s_main));
    CU(cudaStreamSynchronize(e->s_main));
    return SDRX_OK;
}

extern "C" int sdrx_engine_get_outputs(sdrx_engine *e, const sdrx_block_outputs *out) {
    if (!e || !out) return fail(SDRX_ERR_ARG, "null argument");
    if (e->blocks == 0) return fail(SDRX_ERR_STATE, "no block pushed yet");
    if (e->flags & SDRX_FLAG_NO_DEMOD) return fail(SDRX_ERR_STATE, "engine was created with SDRX_FLAG_NO_DEMOD");
    if (!e->last_demod) return fail(SDRX_ERR_STATE, "the last block did not yield 512 samples per channel; boundary B did not run");
    DevGuard g(e->dev);
    int rc = wait_side(e, (int)((e->blocks - 1) % e->ring));
    if (!rc) rc = copy_outputs(e, e->out[(e->blocks - 1) % e->ring], out, e->s_main, true);
    if (rc) return rc;
    CU(cudaStreamSynchronize(e->s_main));
    return SDRX_OK;
}

// Asynchronous form of get_outputs for streaming callers: the D2H copies of the LAST pushed block are queued on
// their own stream behind that block's kernels and the call returns; the next push_block may follow at once, so
// its H2D and kernels overlap this block's read-back.  The (pinned) host buffers are valid after
// sdrx_engine_wait_outputs(e, back) with back = how many blocks have been pushed since.
extern "C" int sdrx_engine_fetch_outputs(sdrx_engine *e, const sdrx_block_outputs *out) {
    if (!e || !out) return fail(SDRX_ERR_ARG, "null argument");
    if (e->blocks == 0) return fail(SDRX_ERR_STATE, "no block pushed yet");
    if (e->flags & SDRX_FLAG_NO_DEMOD) return fail(SDRX_ERR_STATE, "engine was created with SDRX_FLAG_NO_DEMOD");
    if (!e->last_demod) return fail(SDRX_ERR_STATE, "the last block did not yield 512 samples per channel; boundary B did not run");
    DevGuard g(e->dev);
    const int idx = (int)((e->blocks - 1) % e->ring);
    if (e->side_pending[idx]) CU(cudaStreamWaitEvent(e->s_out, e->ev_side_done[idx], 0));
    int rc = copy_outputs(e, e->out[idx], out, e->s_out, false);
    if (rc) return rc;
    CU(cudaEventRecord(e->ev_fetch_done[idx], e->s_out));
    e->fetch_pending[idx] = true;
    return SDRX_OK;
}

extern "C" int sdrx_engine_wait_outputs(sdrx_engine *e, int back) {
    if (!e || back < 0 || back > e->ring - 2) return fail(SDRX_ERR_ARG, "back must be 0 (last pushed block) .. 8");
    if (e->blocks < (uint64_t)back + 1) return fail(SDRX_ERR_STATE, "no such block");
    DevGuard g(e->dev);
    const int idx = (int)((e->blocks - 1 - back) % e->ring);
    if (!e->fetch_pending[idx]) return fail(SDRX_ERR_STATE, "no fetch was queued for that block");
    CU(cudaEventSynchronize(e->ev_fetch_done[idx]));
    return SDRX_OK;
}

extern "C" int sdrx_engine_demod_block(sdrx_engine *e, const float *ch_iq, float pwr_dbfs) {
    if (!e || !ch_iq) return fail(SDRX_ERR_ARG, "null argument");
    if (e->flags & SDRX_FLAG_NO_DEMOD) return fail(SDRX_ERR_STATE, "engine was created with SDRX_FLAG_NO_DEMOD");
    DevGuard g(e->dev);
    const int idx = (int)(e->blocks % e->ring);
    OutSet &o = e->out[idx];
    { int rc = wait_side(e, idx); if (rc) return rc; }
    CU(cudaMemcpy2DAsync(o.ch_iq, (size_t)e->chiq_stride * sizeof(float2), ch_iq, (size_t)kChunk * sizeof(float2),
                         (size_t)kChunk * sizeof(float2), (size_t)e->C, cudaMemcpyHostToDevice, e->s_main));
    CU(cudaMemcpyAsync(o.pwr, &pwr_dbfs, sizeof(float), cudaMemcpyHostToDevice, e->s_main));
    CU(cudaStreamSynchronize(e->s_main));     // pwr_dbfs lives on this stack frame
    int rc = launch_demod(e, idx, false, 0);
    if (rc) return rc;
    e->last_out_len = kChunk;
    e->last_demod = true;
    ++e->blocks;      // takes the next output set like a pushed block; the channelizer (raw ping-pong, e->pushes) is untouched
    return SDRX_OK;
}

extern "C" int sdrx_engine_finish_mix(sdrx_engine *e, const float *parts, int n_parts, float *mix_pre, float *mix, int16_t *pcm) {
    if (!e || !parts || n_parts < 1) return fail(SDRX_ERR_ARG, "null argument");
    if (e->NS > 1) return fail(SDRX_ERR_UNSUPPORTED, "finish_mix: single-stream engines only");
    if (e->flags & SDRX_FLAG_NO_DEMOD) return fail(SDRX_ERR_STATE, "engine was created with SDRX_FLAG_NO_DEMOD");
    DevGuard g(e->dev);
    const size_t bytes = sizeof(float) * 2 * kChunk * (size_t)n_parts;
    if ((size_t)n_parts > e->parts_cap) {
        CU(cudaStreamSynchronize(e->s_main));
        cudaFree(e->d_parts);
        e->d_parts = nullptr;
        CU(cudaMalloc((void **)&e->d_parts, bytes));
        e->parts_cap = (size_t)n_parts;
    }
    const int idx = e->blocks ? (int)((e->blocks - 1) % e->ring) : 0;
    OutSet &o = e->out[idx];
    { int rc = wait_side(e, idx); if (rc) return rc; }
    CU(cudaMemcpyAsync(e->d_parts, parts, bytes, cudaMemcpyHostToDevice, e->s_main));
    { int rc = launch_mix(e, e->s_main, o, nullptr, 0, 0, nullptr, e->d_parts, n_parts, 1); if (rc) return rc; }
    if (mix_pre) CU(cudaMemcpyAsync(mix_pre, o.mix_pre, sizeof(float) * 2 * kChunk, cudaMemcpyDeviceToHost, e->s_main));
    if (mix) CU(cudaMemcpyAsync(mix, o.mix, sizeof(float) * 2 * kChunk, cudaMemcpyDeviceToHost, e->s_main));
    if (pcm) CU(cudaMemcpyAsync(pcm, o.pcm, sizeof(int16_t) * 2 * kChunk, cudaMemcpyDeviceToHost, e->s_main));
    CU(cudaStreamSynchronize(e->s_main));
    return SDRX_OK;
}

// Debug: completion times (ms after the moment the timeline was switched on) of every kernel of the last blocks, as the
// streams reached them: [block][start-of-main, prep, bank, last stage, detector, agc, envelope, lfagc, mix].
extern "C" int sdrx_debug_timeline(sdrx_engine *e, int enable) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    DevGuard g(e->dev);
    if (enable) {
        if (!e->tl_base) {
            CU(cudaEventCreate(&e->tl_base));
            for (auto &row : e->tl_ev) for (int k = 0; k < TL_N; ++k) CU(cudaEventCreate(&row[k]));
        }
        CU(cudaEventRecord(e->tl_base, e->s_main));
        e->tl_first = e->blocks;
    }
    e->tl_on = enable != 0;
    return SDRX_OK;
}
extern "C" int sdrx_debug_timeline_read(sdrx_engine *e, float *out, int max_blocks) {
    if (!e || !out || !e->tl_base) return 0;
    DevGuard g(e->dev);
    if (sdrx_engine_sync(e) != SDRX_OK) return 0;
    const uint64_t n = std::min<uint64_t>({(uint64_t)max_blocks, e->blocks - e->tl_first, (uint64_t)sdrx_engine::kTlBlocks});
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t b = e->blocks - n + i;
        for (int k = 0; k < TL_N; ++k) {
            float ms = -1.0f;
            if (cudaEventElapsedTime(&ms, e->tl_base, e->tl_ev[b % sdrx_engine::kTlBlocks][k]) != cudaSuccess) { cudaGetLastError(); ms = -1.0f; }
            out[i * TL_N + k] = ms;
        }
    }
    return (int)n;
}

extern "C" int sdrx_engine_wait_stream(sdrx_engine *e, void *cuda_stream) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    DevGuard g(e->dev);
    CU(cudaEventRecord(e->ev_ext, (cudaStream_t)cuda_stream));
    CU(cudaStreamWaitEvent(e->s_copy, e->ev_ext, 0));
    return SDRX_OK;
}
extern "C" int sdrx_engine_stream_wait_ingest(sdrx_engine *e, void *cuda_stream) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    if (e->pushes == 0) return SDRX_OK;
    DevGuard g(e->dev);
    CU(cudaStreamWaitEvent((cudaStream_t)cuda_stream, e->ev_copied[(e->pushes + kNBuf - 1) % kNBuf], 0));
    return SDRX_OK;
}
extern "C" int sdrx_engine_stream_wait_outputs(sdrx_engine *e, void *cuda_stream) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    if (e->blocks == 0) return SDRX_OK;
    DevGuard g(e->dev);
    const int idx = (int)((e->blocks - 1) % e->ring);
    if (e->side_pending[idx]) CU(cudaStreamWaitEvent((cudaStream_t)cuda_stream, e->ev_side_done[idx], 0));
    else CU(cudaStreamWaitEvent((cudaStream_t)cuda_stream, e->ev_head_done, 0));
    return SDRX_OK;
}
extern "C" int sdrx_engine_device_outputs(sdrx_engine *e, int back, sdrx_block_outputs *dev) {
    if (!e || !dev || back < 0 || back > e->ring - 2) return fail(SDRX_ERR_ARG, "back must be 0 (last pushed block) .. 8");
    if (e->blocks < (uint64_t)back + 1) return fail(SDRX_ERR_STATE, "no such block");
    const OutSet &o = e->out[(e->blocks - 1 - back) % e->ring];
    memset(dev, 0, sizeof(*dev));
    dev->audio = nullptr;         /* on the device the audio is sample-major: not part of this view */
    dev->sql_state = o.sql; dev->snr = o.snr; dev->mix_pre = o.mix_pre; dev->mix = o.mix; dev->pcm = o.pcm;
    dev->pwr_dbfs = o.pwr; dev->agc_gain = o.agc_gain; dev->lf_gain = o.lf_gain; dev->audio_chan = o.audio_idx; dev->n_audio = o.n_audio;
    return SDRX_OK;
}
extern "C" int sdrx_engine_finish_mix_async(sdrx_engine *e, const float *d_parts, int n_parts, void *cuda_stream, float *mix_pre,
                                            float *mix, int16_t *pcm) {
    if (!e || !d_parts || n_parts < 1) return fail(SDRX_ERR_ARG, "null argument");
    if (e->flags & SDRX_FLAG_NO_DEMOD) return fail(SDRX_ERR_STATE, "engine was created with SDRX_FLAG_NO_DEMOD");
    DevGuard g(e->dev);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int k = (int)(e->n_fin++ % sdrx_engine::kNFin);
    if (!e->d_fin_mix_pre[k]) {
        int rc;
        if ((rc = dalloc(&e->d_fin_mix_pre[k], 2 * kChunk)) || (rc = dalloc(&e->d_fin_mix[k], 2 * kChunk)) ||
            (rc = dalloc(&e->d_fin_pcm[k], 2 * kChunk)))
            return rc;
    }
    OutSet o{};
    o.mix_pre = e->d_fin_mix_pre[k]; o.mix = e->d_fin_mix[k]; o.pcm = e->d_fin_pcm[k];
    { int rc = launch_mix(e, st, o, nullptr, 0, 0, nullptr, d_parts, n_parts, 1); if (rc) return rc; }
    if (mix_pre) CU(cudaMemcpyAsync(mix_pre, o.mix_pre, sizeof(float) * 2 * kChunk, cudaMemcpyDeviceToHost, st));
    if (mix) CU(cudaMemcpyAsync(mix, o.mix, sizeof(float) * 2 * kChunk, cudaMemcpyDeviceToHost, st));
    if (pcm) CU(cudaMemcpyAsync(pcm, o.pcm, sizeof(int16_t) * 2 * kChunk, cudaMemcpyDeviceToHost, st));
    return SDRX_OK;
}

extern "C" void *sdrx_engine_stream(sdrx_engine *e) { return e ? (void *)e->s_main : nullptr; }
extern "C" uint64_t sdrx_engine_launch_count(sdrx_engine *e) { return e ? e->launches : 0; }
// host-only: plans the channel-bank schedule for `groups` slot groups x U units on n_sms SMs and self-checks it
extern "C" int sdrx_debug_bank_plan_check(int groups, int U, int n_sms, int *n_ctas, int *n_pieces, int *max_steps) {
    FastPlan plan;
    if (groups < 1 || U < 1 || n_sms < 1) return -1;
    if (fast_plan_schedule(groups, U, n_sms, plan) != SDRX_OK) return -1;
    if (n_ctas) *n_ctas = (int)plan.ctas.size();
    if (n_pieces) *n_pieces = (int)plan.pieces.size();
    if (max_steps) {                 // steps (125-sample units incl. warm-up) of the busiest CTA: the kernel's critical path
        int worst = 0;
        for (const FastCta &c : plan.ctas) {
            int steps = 0;
            for (int i = c.piece_begin; i < c.piece_end; ++i)
                steps += ((plan.pieces[i].u_begin == 0 && !fast_cut(plan.pieces[i])) ? kFastWarmStart : 1) + plan.pieces[i].seg_len;
            worst = std::max(worst, steps);
        }
        *max_steps = worst;
    }
    return fast_check_plan(groups, U, plan);
}

// Debug: the schedule itself, one record of 6 ints per piece: cta, group, u_begin, u_end, seg_len, steps.
extern "C" int sdrx_debug_bank_plan_dump(int groups, int U, int n_sms, int *out, int max_records) {
    FastPlan plan;
    if (groups < 1 || U < 1 || n_sms < 1 || !out) return -1;
    if (fast_plan_schedule(groups, U, n_sms, plan) != SDRX_OK) return -1;
    int n = 0;
    for (size_t c = 0; c < plan.ctas.size(); ++c)
        for (int i = plan.ctas[c].piece_begin; i < plan.ctas[c].piece_end && n < max_records; ++i, ++n) {
            const FastPiece &p = plan.pieces[i];
            int *r = out + 6 * n;
            r[0] = (int)c; r[1] = p.group; r[2] = p.u_begin; r[3] = p.u_end; r[4] = p.seg_len; r[5] = fast_piece_steps(p);
        }
    return n;
}

// Debug: per-warp start / end times (globaltimer ns) of the next channel-bank launches.  enable: allocates
// [max_ctas][16][2] u64 and points the kernel at it; read: copies it out and returns the number of u64 written.
static unsigned long long *g_trace_buf = nullptr;
static const int kTraceCtas = 160;
extern "C" int sdrx_debug_bank_trace(int enable) {
    unsigned long long *p = nullptr;
    if (enable) {
        if (!g_trace_buf) CU(cudaMalloc((void **)&g_trace_buf, sizeof(unsigned long long) * kTraceCtas * 16 * 2));
        CU(cudaMemset(g_trace_buf, 0, sizeof(unsigned long long) * kTraceCtas * 16 * 2));
        p = g_trace_buf;
    }
    CU(cudaMemcpyToSymbol(g_bank_trace, &p, sizeof(p)));
    return SDRX_OK;
}
extern "C" int sdrx_debug_bank_trace_read(unsigned long long *out, int max_u64) {
    if (!g_trace_buf || !out) return 0;
    const int n = std::min(max_u64, kTraceCtas * 16 * 2);
    if (cudaMemcpy(out, g_trace_buf, sizeof(unsigned long long) * n, cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
    return n;
}

extern "C" int sdrx_engine_uses_fast_path(sdrx_engine *e) { return e && (e->use_fast || e->use_tgemm) ? 1 : 0; }
extern "C" int sdrx_engine_bank_pairs(sdrx_engine *e) { return !e ? 0 : (e->use_tgemm ? e->tg.n_pairs : (e->use_fast ? e->fast.n_pairs : 0)); }
extern "C" int sdrx_engine_uses_table_gemm(sdrx_engine *e) { return e && e->use_tgemm ? 1 : 0; }
extern "C" int sdrx_engine_uses_tensor_cores(sdrx_engine *e) { return e && e->use_tgemm && e->use_tc ? 1 : 0; }
extern "C" int sdrx_engine_set_timing(sdrx_engine *e, int enable) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    e->timing = enable != 0;
    e->timing_valid = false;
    return SDRX_OK;
}
extern "C" int sdrx_engine_last_timing(sdrx_engine *e, float *bank_ms, float *block_ms) {
    if (!e) return fail(SDRX_ERR_ARG, "null engine");
    if (!e->timing_valid) return fail(SDRX_ERR_STATE, "timing not enabled or no block pushed since");
    DevGuard g(e->dev);
    CU(cudaEventSynchronize(e->ev_t2));
    float a = 0, b = 0;
    if (e->bank_timed) CU(cudaEventElapsedTime(&a, e->ev_b0, e->ev_b1));   // the specialised bank kernel alone
    else CU(cudaEventElapsedTime(&a, e->ev_t0, e->ev_t1));
    CU(cudaEventElapsedTime(&b, e->ev_t0, e->ev_t2));
    if (bank_ms) *bank_ms = a;
    if (block_ms) *block_ms = b;
    return SDRX_OK;
}

// ------------------------------------------------------------------------------------------------
// Carried state.  Layout: StateHeader, raw history (raw_hist * 8 bytes, wire format `fmt`), then for
// each stage s >= 1 its [C][H-1] delay line, then AGC gain/power, LfAGC gain/power (float[C] each),
// sql_state, sql_state_prev (u8[C] each), the FM demodulators' previous samples (float2[C]) and the FIR2 ring as
// (ntaps-1) stereo frames, oldest first.
namespace {

struct StateHeader {
    uint32_t magic, version;
    int32_t  C, S, fmt, raw_hist, fir2_ntaps, demod;
    uint64_t total;
    int64_t  trans_pos;
    uint64_t blocks;
    int32_t  phase[SDRX_MAX_STAGES];
    int32_t  H[SDRX_MAX_STAGES];
    int32_t  ft_k, ftfir;
};

size_t state_bytes(const sdrx_engine *e) {
    size_t n = sizeof(StateHeader) + (size_t)e->raw_hist * 8;
    for (int s = 1; s < e->S; ++s) n += sizeof(float2) * (size_t)(e->H[s] - 1) * e->C;
    if (!(e->flags & SDRX_FLAG_NO_DEMOD))
        n += (size_t)e->C * (4 * sizeof(float) + 2 + sizeof(float2)) + sizeof(float) * 2 * (size_t)(e->fir2_ntaps - 1);
    return n;
}

}  // namespace

extern "C" size_t sdrx_engine_state_size(sdrx_engine *e) { return e ? state_bytes(e) : 0; }

extern "C" int sdrx_engine_save_state(sdrx_engine *e, void *buf, size_t size) {
    if (!e || !buf) return fail(SDRX_ERR_ARG, "null argument");
    if (e->NS > 1) return fail(SDRX_ERR_UNSUPPORTED, "state save / load of a multi-stream engine is not implemented");
    if (e->poly) return fail(SDRX_ERR_UNSUPPORTED, "state save / load of a polyphase engine is not implemented");
    if (size < state_bytes(e)) return fail(SDRX_ERR_ARG, "state buffer too small");
    DevGuard g(e->dev);
    CU(cudaStreamSynchronize(e->s_copy));
    CU(cudaStreamSynchronize(e->s_prep));
    CU(cudaStreamSynchronize(e->s_main));
    CU(cudaStreamSynchronize(e->s_side));
    CU(cudaStreamSynchronize(e->s_recur));
    CU(cudaStreamSynchronize(e->s_lf));
    CU(cudaStreamSynchronize(e->s_mix));
    if (e->generic_stale) {      // the blob always carries the stage delay lines (a generic engine can load it)
        const int last_ = (int)((e->pushes + kNBuf - 1) % kNBuf);
        int rc = materialize_generic(e, e->fmt, e->d_raw[last_], (long)e->raw_hist + (long)e->prev_n, e->total);
        if (rc) return rc;
        CU(cudaStreamSynchronize(e->s_main));
    }
    uint8_t *p = (uint8_t *)buf;
    StateHeader h{};
    h.magic = kStateMagic; h.version = 3;
    h.C = e->C; h.S = e->S; h.fmt = e->fmt; h.raw_hist = e->raw_hist; h.total = e->total; h.fir2_ntaps = e->fir2_ntaps;
    h.demod = (e->flags & SDRX_FLAG_NO_DEMOD) ? 0 : 1;
    h.trans_pos = e->trans_pos; h.blocks = e->pushes;
    for (int s = 0; s < SDRX_MAX_STAGES; ++s) { h.phase[s] = e->phase[s]; h.H[s] = e->H[s]; }
    h.ft_k = e->ft_k; h.ftfir = e->ftfir ? 1 : 0;
    memcpy(p, &h, sizeof(h)); p += sizeof(h);
    const int bps = e->fmt >= 0 ? raw_bytes_per_sample(e->fmt) : 8;
    const int last = (int)((e->pushes + kNBuf - 1) % kNBuf);
    memset(p, 0, (size_t)e->raw_hist * 8);
    CU(cudaMemcpy(p, e->d_raw[last] + (size_t)e->prev_n * bps, (size_t)e->raw_hist * bps, cudaMemcpyDeviceToHost));
    p += (size_t)e->raw_hist * 8;
    for (int s = 1; s < e->S; ++s) {
        const size_t row = sizeof(float2) * (size_t)(e->H[s] - 1);
        if (row) CU(cudaMemcpy2D(p, row, e->d_stage_in[s], sizeof(float2) * (size_t)e->stage_stride[s], row, e->C, cudaMemcpyDeviceToHost));
        p += row * e->C;
    }
    if (h.demod) {
        const size_t C = (size_t)e->C;
        CU(cudaMemcpy(p, e->st.agc_gain, C * 4, cudaMemcpyDeviceToHost)); p += C * 4;
        CU(cudaMemcpy(p, e->st.agc_power, C * 4, cudaMemcpyDeviceToHost)); p += C * 4;
        CU(cudaMemcpy(p, e->st.lf_gain, C * 4, cudaMemcpyDeviceToHost)); p += C * 4;
        CU(cudaMemcpy(p, e->st.lf_power, C * 4, cudaMemcpyDeviceToHost)); p += C * 4;
        CU(cudaMemcpy(p, e->st.sql_state, C, cudaMemcpyDeviceToHost)); p += C;
        CU(cudaMemcpy(p, e->st.sql_prev, C, cudaMemcpyDeviceToHost)); p += C;
        CU(cudaMemcpy(p, e->d_fm[e->fm_cur], C * sizeof(float2), cudaMemcpyDeviceToHost)); p += C * sizeof(float2);
        const size_t fb = sizeof(float) * 2 * (size_t)(e->fir2_ntaps - 1);
        if (fb) CU(cudaMemcpy(p, e->d_fir_hist, fb, cudaMemcpyDeviceToHost));
        p += fb;
    }
    return SDRX_OK;
}

extern "C" int sdrx_engine_load_state(sdrx_engine *e, const void *buf, size_t size) {
    if (!e || !buf) return fail(SDRX_ERR_ARG, "null argument");
    if (e->NS > 1) return fail(SDRX_ERR_UNSUPPORTED, "state save / load of a multi-stream engine is not implemented");
    if (size < sizeof(StateHeader)) return fail(SDRX_ERR_ARG, "state buffer too small");
    StateHeader h;
    memcpy(&h, buf, sizeof(h));
    if (h.magic != kStateMagic || h.version != 3) return fail(SDRX_ERR_ARG, "not an sdrx_b200 state blob");
    if (h.C != e->C || h.S != e->S || h.raw_hist != e->raw_hist || h.fir2_ntaps != e->fir2_ntaps ||
        h.demod != ((e->flags & SDRX_FLAG_NO_DEMOD) ? 0 : 1) || h.ftfir != (e->ftfir ? 1 : 0))
        return fail(SDRX_ERR_ARG, "state blob was saved by a differently configured engine");
    for (int s = 0; s < e->S; ++s)
        if (h.H[s] != e->H[s]) return fail(SDRX_ERR_ARG, "state blob stage layout differs");
    if (size < state_bytes(e)) return fail(SDRX_ERR_ARG, "state buffer too small");
    DevGuard g(e->dev);
    CU(cudaStreamSynchronize(e->s_copy));
    CU(cudaStreamSynchronize(e->s_prep));
    CU(cudaStreamSynchronize(e->s_main));
    CU(cudaStreamSynchronize(e->s_side));
    CU(cudaStreamSynchronize(e->s_recur));
    CU(cudaStreamSynchronize(e->s_lf));
    CU(cudaStreamSynchronize(e->s_mix));
    CU(cudaStreamSynchronize(e->s_out));
    for (int b = 0; b < kNOut; ++b) e->side_pending[b] = e->fetch_pending[b] = false;
    e->tail_pending[0] = e->tail_pending[1] = false;
    e->last_tail_par = -1;
    const uint8_t *p = (const uint8_t *)buf + sizeof(h);
    e->fmt = h.fmt; e->trans_pos = h.trans_pos; e->blocks = e->pushes = h.blocks; e->total = h.total; e->generic_stale = false;
    for (int s = 0; s < SDRX_MAX_STAGES; ++s) e->phase[s] = h.phase[s];
    e->ft_k = h.ft_k;
    e->prev_n = 0;          // the history is parked at the very front of the "previous" buffer
    e->last_out_len = 0; e->last_demod = false; e->timing_valid = false;
    const int last = (int)((e->pushes + kNBuf - 1) % kNBuf);
    CU(cudaMemcpy(e->d_raw[last], p, (size_t)e->raw_hist * 8, cudaMemcpyHostToDevice));
    p += (size_t)e->raw_hist * 8;
    for (int s = 1; s < e->S; ++s) {
        const size_t row = sizeof(float2) * (size_t)(e->H[s] - 1);
        if (row) CU(cudaMemcpy2D(e->d_stage_in[s], sizeof(float2) * (size_t)e->stage_stride[s], p, row, row, e->C, cudaMemcpyHostToDevice));
        p += row * e->C;
    }
    if (h.demod) {
        const size_t C = (size_t)e->C;
        CU(cudaMemcpy(e->st.agc_gain, p, C * 4, cudaMemcpyHostToDevice)); p += C * 4;
        CU(cudaMemcpy(e->st.agc_power, p, C * 4, cudaMemcpyHostToDevice)); p += C * 4;
        CU(cudaMemcpy(e->st.lf_gain, p, C * 4, cudaMemcpyHostToDevice)); p += C * 4;
        CU(cudaMemcpy(e->st.lf_power, p, C * 4, cudaMemcpyHostToDevice)); p += C * 4;
        CU(cudaMemcpy(e->st.sql_state, p, C, cudaMemcpyHostToDevice)); p += C;
        CU(cudaMemcpy(e->st.sql_prev, p, C, cudaMemcpyHostToDevice)); p += C;
        CU(cudaMemcpy(e->d_fm[e->fm_cur], p, C * sizeof(float2), cudaMemcpyHostToDevice)); p += C * sizeof(float2);
        const size_t fb = sizeof(float) * 2 * (size_t)(e->fir2_ntaps - 1);
        if (fb) CU(cudaMemcpy(e->d_fir_hist, p, fb, cudaMemcpyHostToDevice));
        p += fb;
    }
    return SDRX_OK;
}

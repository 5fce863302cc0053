/*
 * sdrx_b200 -- C ABI of the B200-native sdrx per-channel DSP chain.
 *
 * This is the drop-in boundary.  Plain pointers and sizes only; the caller owns
 * every host buffer, the engine owns all device memory.  Citations are file:line under
 * /root/reference/src of the interface each entry point replaces.
 *
 *   Boundary A (channelizer)   replaces sdrx.cpp:327-346  (the per-channel msd.decimate loop of
 *                              data_cb) and, one channel at a time, MSD::decimate msd.hpp:75
 *                              and DS::addJob ds.hpp:43.
 *   Boundary B (demod/squelch) replaces sdrx.cpp:452-573 (AGC, AM envelope, LfAGC, ramps, pan
 *                              mix, window+FFT, SNR squelch) and :646-656 (FIR2, float->s16).
 *   Ingest                     replaces rtl_dev.cpp:316-341 / airspy_dev.cpp:355-405 (raw->float
 *                              and block power), fused into the first kernel.
 *
 * Error convention: the reference's DSP classes cannot fail (void / value returns,
 * SURVEY.md §8b "Error conventions"); every function here that can hit a CUDA error returns int:
 * 0 on success, a negative sdrx_status otherwise, and sdrx_last_error() gives the text.  There is
 * no CPU fallback: without a usable CUDA device creation fails with SDRX_ERR_CUDA.
 *
 * Threading: one engine per GPU, not re-entrant per engine (the reference calls side A from one
 * device thread and side B from the ALSA thread; here both sides of one block are queued on the
 * engine's CUDA streams by whichever host thread calls).
 */
#ifndef SDRX_B200_H
#define SDRX_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SDRX_CH_IQ_BUF_SIZE 512   /* sdrx.cpp:129 CH_IQ_BUF_SIZE == FFT_SIZE */
#define SDRX_MAX_STAGES 8

typedef enum {
    SDRX_OK = 0,
    SDRX_ERR_ARG = -1,        /* bad argument / configuration                                  */
    SDRX_ERR_CUDA = -2,       /* CUDA runtime error or no device                               */
    SDRX_ERR_STATE = -3,      /* call out of order (e.g. get before push)                      */
    SDRX_ERR_UNSUPPORTED = -4 /* configuration outside what the engine implements              */
} sdrx_status;

/* Wire formats of one IQ block (interleaved I,Q).
 *   U8   rtl_sdr / librtlsdr: float = u8/127.5f - 1.0f          rtl_dev.cpp:320-323
 *   S16  airspy INT16_IQ:     float = s16 * (1/32768)           libairspy airspy.c:58-62,:304
 *   F32  airspy FLOAT32_IQ == iqsample_t                        airspy_dev.cpp:355-357           */
typedef enum { SDRX_FMT_U8 = 0, SDRX_FMT_S16 = 1, SDRX_FMT_F32 = 2 } sdrx_iq_format;

/* flags for sdrx_engine_config.flags */
#define SDRX_FLAG_FORCE_GENERIC 0x1u  /* never take the specialised channel-bank kernel         */
#define SDRX_FLAG_NO_DEMOD      0x2u  /* channelizer only (boundary A); skip boundary B kernels  */
#define SDRX_FLAG_NO_PAIRING    0x8u  /* specialised bank: one channel per lane slot even where +-offset channels
                                         could share a translator table as conjugates                       */
#define SDRX_FLAG_FTFIR         0x10u /* --ftfir (sdrx.cpp:1199,:1692): the first stage is the frequency-translating FIR,
                                         msd.hpp:104-131 + calculateOutputTranslated :341-423, instead of mix + FIR      */
#define SDRX_FLAG_POLYPHASE     0x20u /* OPT-IN, NOT BIT-COMPATIBLE: boundary A as a polyphase / DFT channelizer (the four MSD stages as one
                                         8857-tap prototype folded into a 1200-point DFT per output instant -- what the reference's
                                         README.md:25-28 considers for many simultaneous channels).  Same map as the chain in exact
                                         arithmetic, ~100x fewer operations; differs from the reference by its float-rounded translator
                                         table (sdrx.cpp:1682-1690).  Needs the 10 MS/s chain (N = 1200), ch_offset, whole blocks.      */
#define SDRX_FLAG_TGEMM         0x40u /* OPT-IN: boundary A as fold + table GEMM (tgemm.cuh) -- the four MSD stages as one prototype folded
                                         modulo the translator period, then Y = T U with the reference's OWN translator tables as the
                                         matrix (the "complex GEMM (channels x taps)" recast).  Same map as the chain; differs from it by
                                         FP32 rounding only, but its rounding is independent of the reference's: audio and squelch
                                         decisions meet the parity bar, the channel IQ of closed, noise-only channels next to strong
                                         carriers reaches 1.4e-4 of relative RMS (time-domain bank: 4e-5).  Any chain whose translator
                                         period is a multiple of 16; whole blocks.                                                     */
#define SDRX_FLAG_TGEMM_TC      0x80u /* OPT-IN: SDRX_FLAG_TGEMM with the matrix product on the 5th-generation tensor cores (tgemm_tc.cuh:
                                         tcgen05.mma, accumulators in tensor memory, operands through TMA).  Every FP32 operand is split
                                         into three bf16 planes and all nine plane products are kept, so the products are exact; partial
                                         sums over 16 table entries are added in FP32 (round to nearest) as in the FP32 kernel.  Same bars
                                         as SDRX_FLAG_TGEMM.                                                                            */
#define SDRX_FLAG_PARTIAL_MIX   0x4u  /* channel-sharded run: stop at mix_pre (this GPU's share of the pan
                                         mix); FIR2 + s16 run once, in sdrx_engine_finish_mix              */

#define SDRX_MOD_AM 0       /* Modulation::AM (sdrx.cpp:151) */
#define SDRX_MOD_FM 1       /* Modulation::FM */

typedef struct sdrx_engine sdrx_engine;

/* Everything sdrx.cpp:1584-1714 sets up per run.  Arrays are copied at creation. */
typedef struct {
    int            device;          /* CUDA device ordinal                                        */
    int            n_channels;      /* settings.channels.size()                                   */
    int            translator_len;  /* N (sdrx.cpp:1571-1583)                                     */
    const float   *translators;     /* [C][N][2] re,im; row ignored where has_translator[c]==0     */
    const uint8_t *has_translator;  /* [C]; 0 == empty translator vector (ch_offset==0, :1685)     */
    int            n_stages;        /* MSD::Stage list (msd.hpp:46-49)                             */
    const int     *m;               /* [n_stages] down-sampling factors                            */
    const int     *ntaps;           /* [n_stages] (odd, symmetric taps; msd.hpp:169-170)           */
    const float   *taps;            /* concatenated                                                */
    unsigned       block_len;       /* largest block (complex samples) push_block will be given   */
    const float   *sql_level;       /* [C] dB, Channel::sql_level (sdrx.cpp:215); NULL -> 9.0      */
    const int     *pos;             /* [C] audio position -2..2 (sdrx.cpp:1024-1044); NULL -> 0    */
    /* AGC / LfAGC run parameters.  Zero-initialised fields take the values sdrx.cpp:1702-1710 sets:
     * reference 1, attack 1, decay 0.01, max gain 300 (AGC) / 200 (LfAGC), alpha .01 / .005; the
     * detectors start from the constructors' reference (0.25, agc.hpp:30,:86).                   */
    float          agc_attack, agc_decay, agc_reference, agc_max_gain, agc_alpha;
    float          lf_attack, lf_decay, lf_reference, lf_max_gain, lf_alpha;
    int            use_lf_agc;      /* --lf-agc (sdrx.cpp:1710)                                    */
    float          lf_gain_db;      /* --volume: FIR2::setGain (sdrx.cpp:881)                      */
    const float   *fir2_taps;       /* coeff_bp4am_channel (coeffs.hpp:1812)                       */
    int            fir2_ntaps;
    const float   *passband_shape;  /* [512] coeffs.hpp:2018                                       */
    const float   *ramp_up;         /* [512] coeffs.hpp:2541                                       */
    const float   *ramp_down;       /* [512] coeffs.hpp:3064                                       */
    unsigned       flags;
    const uint8_t *modulation;      /* [C] SDRX_MOD_AM / SDRX_MOD_FM: Channel::mod (sdrx.cpp:230, the
                                       NNN.NNN/SQL/AM|FM channel syntax :1351-1358, --modulation :1197);
                                       NULL -> AM.  FM is Demod::demod's discriminator (sdrx.cpp:181-194). */
    const int     *ch_offset;       /* [C] channel offsets in 8.33 kHz steps from the tuner (channel_to_offset, sdrx.cpp:1449-1478): only
                                       SDRX_FLAG_POLYPHASE needs them (the DFT bin of a channel); NULL otherwise                    */
    int            n_streams;       /* 0 / 1: one IQ stream (one dongle, sdrx.cpp:1742-1745).  > 1: that many independent
                                       streams, all listening to the SAME channel plan (the arrays above describe one
                                       stream), processed in lockstep by ONE launch sequence per block (BASELINE.json config 4:
                                       many small receivers on one GPU).  push_block then takes n_streams blocks back to
                                       back ([n_streams][n_samples]); every per-channel result has n_streams * n_channels
                                       entries, stream-major; mix_pre / mix / pcm / pwr_dbfs / n_audio have one entry (or 1024)
                                       per stream, and audio_chan holds stream-local channel numbers, n_channels per stream.
                                       Needs the specialised bank (10 MS/s chain), whole blocks, n_channels % 16 == 0.       */
} sdrx_engine_config;

/* Per-block outputs of boundary B.  Any pointer may be NULL. */
typedef struct {
    float   *audio;      /* [C][512]  the sample that is mixed per channel (0 where none)          */
    uint8_t *sql_state;  /* [C]       squelch state after this block's decision (sdrx.cpp:569-573)  */
    float   *snr;        /* [C]       sdrx.cpp:566                                                  */
    /* (multi-stream engines: C = n_streams * n_channels below, [1024] and [1] are per stream, stream-major) */
    float   *mix_pre;    /* [1024]    stereo mix before FIR2 (sdrx.cpp:471-532)                     */
    float   *mix;        /* [1024]    after FIR2 (sdrx.cpp:646)                                     */
    int16_t *pcm;        /* [1024]    sdrx.cpp:649-656                                              */
    float   *pwr_dbfs;   /* [1]       Metadata.pwr_dbfs (rtl_dev.cpp:329-341)                       */
    float   *agc_gain;   /* [C]       AGC::gain()   (agc.hpp:46)                                    */
    float   *lf_gain;    /* [C]       LfAGC::gain() (agc.hpp:101)                                   */
    /* Sparse form of `audio`: only channels whose squelch is open or ramping down contribute to the mix
     * (sdrx.cpp:459-463, :498-501); every other row of `audio` is exact zeros.  n_audio[0] = number of such channels,
     * audio_chan[k] = their indices (ascending), audio_rows[k][512] = their audio, k < n_audio[0]; later rows are left
     * untouched.  With sdrx_engine_fetch_outputs these three must be pinned (sdrx_host_alloc): the device writes them
     * straight across the host link, so a block costs 2 KB per OPEN channel instead of 2 KB per channel.         */
    float   *audio_rows; /* [C][512]  capacity; rows [0, n_audio) valid                              */
    int32_t *audio_chan; /* [C]                                                                     */
    int32_t *n_audio;    /* [1]                                                                     */
} sdrx_block_outputs;

const char *sdrx_last_error(void);
const char *sdrx_version(void);
int  sdrx_device_count(void);

int  sdrx_engine_create(const sdrx_engine_config *cfg, sdrx_engine **out);
void sdrx_engine_destroy(sdrx_engine *e);

/* One 32 ms block through the whole chain (data_cb, sdrx.cpp:306 + alsa_write_cb, :420), queued
 * asynchronously.  `raw` is a HOST pointer in `fmt`; it is staged to the device with
 * cudaMemcpyAsync from pinned memory (directly when `raw` came from sdrx_host_alloc, through the
 * engine's own pinned double buffer otherwise) and may be reused as soon as the call returns
 * unless it is pinned, in which case after sdrx_engine_sync / any get.  n_samples may be any
 * length <= block_len for the channelizer; boundary B runs only when the chain yielded exactly
 * 512 output samples per channel.                                                               */
int  sdrx_engine_push_block(sdrx_engine *e, const void *raw, int fmt, unsigned n_samples);
/* Same with the block already resident in device memory (`d_raw` a device pointer). */
int  sdrx_engine_push_block_device(sdrx_engine *e, const void *d_raw, int fmt, unsigned n_samples);
/* Number of channel IQ samples per channel the last pushed block produced. */
int  sdrx_engine_last_out_len(sdrx_engine *e, unsigned *out_len);
/* Boundary A result: the CRB chunk iqsample_t[C][512] (sdrx.cpp:343-344). Waits for the block. */
int  sdrx_engine_get_channel_iq(sdrx_engine *e, float *ch_iq);
/* Boundary B results. Waits for the block. */
int  sdrx_engine_get_outputs(sdrx_engine *e, const sdrx_block_outputs *out);
/* Streaming form: queue the read-back of the last pushed block and return; push the next block(s) at once (their
 * H2D and kernels overlap this read-back); the host buffers (pinned) are valid after
 * sdrx_engine_wait_outputs(e, back), back = number of blocks pushed since the fetch (0..8: the engine keeps a ring
 * of ten output sets, so a caller may run a few blocks ahead of the results it reads -- enough to take the
 * side-stream kernels + device->host->device round trip off the critical path).                                                   */
int  sdrx_engine_fetch_outputs(sdrx_engine *e, const sdrx_block_outputs *out);
int  sdrx_engine_wait_outputs(sdrx_engine *e, int back);
/* Boundary B on caller-supplied channel IQ (host, [C][512][2]) instead of the channelizer's. */
int  sdrx_engine_demod_block(sdrx_engine *e, const float *ch_iq, float pwr_dbfs);
int  sdrx_engine_sync(sdrx_engine *e);
/* The recurrences and the mix of a block run on a side stream (they overlap the next block's channel bank).
 * sdrx_engine_join orders sdrx_engine_stream() behind everything queued there so far, so that an event a caller
 * records on that stream afterwards covers all of every pushed block.  get_* / sync do this implicitly.     */
int  sdrx_engine_join(sdrx_engine *e);
/* Channel-sharded runs: the stereo mix is a sum over channels (sdrx.cpp:471-532), so each GPU
 * returns its partial mix_pre[1024] and ONE engine finishes the block: adds the n_parts partials (host,
 * [n_parts][1024]) in the order given, then FIR2 (sdrx.cpp:646) with this engine's carried ring and float->s16
 * (sdrx.cpp:649-656).  mix_pre / mix / pcm (host, 1024 each) may be NULL.                                  */
int  sdrx_engine_finish_mix(sdrx_engine *e, const float *parts, int n_parts, float *mix_pre, float *mix, int16_t *pcm);

/* ---- plumbing for callers that drive the engine from their own CUDA streams (the channel-sharded receiver,
 * sdrx_b200/sharded.py: NCCL broadcast of the block in, NCCL gather of the partial mixes out) -- no host synchronisation:
 *   wait_stream          the next push_block_device's ingest is ordered behind everything queued on `cuda_stream` so far
 *   stream_wait_ingest   `cuda_stream` waits until the last pushed block has been copied out of the caller's buffer
 *   stream_wait_outputs  `cuda_stream` waits until boundary B of the last pushed block is complete
 *   device_outputs       DEVICE pointers of the results of the block pushed `back` pushes ago (0 = last; valid until eight
 *                        more blocks have been pushed; audio_rows is NULL, audio_chan / n_audio are the device-side list)
 *   finish_mix_async     sdrx_engine_finish_mix with the partials already on the device (`d_parts`, [n_parts][1024]), queued
 *                        on `cuda_stream`; mix_pre / mix / pcm (pinned host, may be NULL) are valid once the caller has
 *                        synchronised with that stream.                                                              */
int  sdrx_engine_wait_stream(sdrx_engine *e, void *cuda_stream);
int  sdrx_engine_stream_wait_ingest(sdrx_engine *e, void *cuda_stream);
int  sdrx_engine_stream_wait_outputs(sdrx_engine *e, void *cuda_stream);
int  sdrx_engine_device_outputs(sdrx_engine *e, int back, sdrx_block_outputs *dev);
int  sdrx_engine_finish_mix_async(sdrx_engine *e, const float *d_parts, int n_parts, void *cuda_stream, float *mix_pre,
                                  float *mix, int16_t *pcm);

/* Carried state (MSD delay lines + trans_pos, AGC/LfAGC gain+power, squelch states, FIR2 ring):
 * what the reference keeps in its objects between blocks (SURVEY.md §5 checkpoint/resume). */
size_t sdrx_engine_state_size(sdrx_engine *e);
int  sdrx_engine_save_state(sdrx_engine *e, void *buf, size_t size);
int  sdrx_engine_load_state(sdrx_engine *e, const void *buf, size_t size);
int  sdrx_engine_reset(sdrx_engine *e);

/* Introspection for the bench / tests. */
void    *sdrx_engine_stream(sdrx_engine *e);          /* cudaStream_t the hot kernels run on       */
uint64_t sdrx_engine_launch_count(sdrx_engine *e);    /* kernels launched so far                   */
int      sdrx_engine_uses_fast_path(sdrx_engine *e);  /* 1 when the specialised channel bank runs  */
int      sdrx_engine_bank_pairs(sdrx_engine *e);      /* conjugate channel pairs sharing a lane slot */
int      sdrx_engine_uses_table_gemm(sdrx_engine *e); /* 1 when boundary A runs as fold + table GEMM   */
int      sdrx_engine_uses_tensor_cores(sdrx_engine *e); /* 1 when that GEMM runs on tcgen05 (SDRX_FLAG_TGEMM_TC) */
/* Device-timed duration (ms, CUDA events on the engine stream) of the channel-bank kernel(s) and of
 * the whole block for the last pushed block; valid after a sync. */
int      sdrx_engine_last_timing(sdrx_engine *e, float *bank_ms, float *block_ms);
int      sdrx_engine_set_timing(sdrx_engine *e, int enable);

/* Host-only self-check of the channel-bank work schedule (no GPU needed): plans `groups` slot groups x U units of
 * 125 samples over n_sms SMs; returns the number of tiling violations (0 = every output produced exactly once).  */
int      sdrx_debug_bank_plan_check(int groups, int U, int n_sms, int *n_ctas, int *n_pieces, int *max_steps);
/* The schedule itself: 6 ints per piece (cta, group, u_begin, u_end, seg_len, steps); returns the piece count. */
int      sdrx_debug_bank_plan_dump(int groups, int U, int n_sms, int *out, int max_records);
/* Debug trace of the channel bank: per-warp start / end times (globaltimer, ns) of the launches that follow;
 * read returns the number of values copied, laid out [cta][16 warps][start, end] (tools/diag_bank_trace.py). */
int      sdrx_debug_bank_trace(int enable);
int      sdrx_debug_bank_trace_read(unsigned long long *out, int max_u64);
/* Per-kernel completion times of the engine's last blocks, as its streams reached them: enable, push >= max_blocks blocks, read.
 * out[block][9] = ms after the enable: start of the block on the main stream, prep, channel bank, last MSD stage, detector, AGC,
 * envelope, LfAGC, mix (-1: not run).  Returns the number of blocks written (<= 16).  tools/diag_timeline.py.            */
int      sdrx_debug_timeline(sdrx_engine *e, int enable);
int      sdrx_debug_timeline_read(sdrx_engine *e, float *out, int max_blocks);
/* Same for the front end's five passes: [pass][CTA-0 start, latest CTA end] of the last call (10 values). */
int      sdrx_debug_fe_trace(int enable);
int      sdrx_debug_fe_trace_read(unsigned long long *out);

/* Pinned host memory for zero-staging push_block / get_*.  */
void *sdrx_host_alloc(size_t bytes);
void  sdrx_host_free(void *p);
/* Device memory helpers for callers that keep IQ resident (bench `value` leg, file replay). */
void *sdrx_device_alloc(int device, size_t bytes);
void  sdrx_device_free(int device, void *p);
int   sdrx_device_upload(int device, void *d_dst, const void *h_src, size_t bytes);

/* Set-up helpers that must match the reference bit for bit and therefore run on the host with the
 * same libm: translator table sdrx.cpp:1682-1690, Hamming window sdrx.cpp:898-900. */
void sdrx_make_translator(int N, int z, int ch_offset, float *out /* [N][2] */);
void sdrx_make_window(float *out /* [512] */);

/* ---- single-object surface behind the drop-in C++ classes (msd.hpp / agc.hpp / fir.hpp) ---- */
typedef struct sdrx_msd sdrx_msd;
int    sdrx_msd_create(const float *translator, int N, int n_stages, const int *m, const int *ntaps,
                       const float *taps, int use_ftfir, int device, sdrx_msd **out);   /* msd.hpp:54 */
int    sdrx_msd_clone(const sdrx_msd *src, sdrx_msd **out);                             /* MSD copy   */
void   sdrx_msd_destroy(sdrx_msd *m);
size_t sdrx_msd_m(const sdrx_msd *m);                                                   /* msd.hpp:71 */
int    sdrx_msd_decimate(sdrx_msd *m, const float *in, size_t in_len, float *out, size_t *out_len); /* :75 */

typedef struct sdrx_agc_state {   /* agc.hpp:72-79 / :134-142 */
    float attack, decay, reference, min_gain, max_gain, alpha, gain, power;
    int   active;                 /* LfAGC only */
} sdrx_agc_state;
int sdrx_agc_adjust(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out);    /* agc.hpp:48  */
int sdrx_lfagc_adjust(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out);  /* agc.hpp:106 */
/* sdrx_agc_adjust for callers that may not allocate (tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in
 * the reference.  Same arguments, same results bit for bit, but the staged input,
 * output and (gain, power) live in one fixed device array inside the library (129 KB on every GPU that loads it, until exit)
 * instead of grow-only cudaMalloc'ed buffers: the call allocates nothing, on the device or through any library.  At most
 * SDRX_AGC_NOALLOC_MAX complex samples per call (SDRX_ERR_ARG beyond that).  Every call on a device shares the array: calls are
 * serialised by a mutex and synchronous, so they never overlap on the GPU through this entry point; a caller that reaches the
 * same device from another process image or bypasses the mutex would race.                                                   */
#define SDRX_AGC_NOALLOC_MAX 8192u
int sdrx_agc_adjust_noalloc(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out);

typedef struct sdrx_fir sdrx_fir;
/* kind 1 FIR (fir.hpp:78), 2 FIR2 (fir.hpp:130), 3 FIR3<iqsample_t> (fir.hpp:28) */
int  sdrx_fir_create(int kind, const float *c, int n, int device, sdrx_fir **out);
int  sdrx_fir_clone(const sdrx_fir *src, sdrx_fir **out);   /* implicit copy of FIR/FIR2/FIR3: taps, gain AND delay line
                                                              (buf_, pos_ are plain members, fir.hpp:69-74, :176-182) */
void sdrx_fir_destroy(sdrx_fir *f);
int  sdrx_fir_set_gain(sdrx_fir *f, float gain_db);                                     /* fir.hpp:51  */
int  sdrx_fir_filter(sdrx_fir *f, const float *in, unsigned in_len, float *out);        /* fir.hpp:34  */

/* ---- Airspy real -> IQ front end: what libairspy does on the host for
 * AIRSPY_SAMPLE_FLOAT32_IQ -- unpack_samples (airspy.c:323-338), convert_samples_float (:311-321) and
 * iqconverter_float_process (iqconverter_float.c:504-508: remove_dc, translate_fs_4, half-band FIR, delay) --
 * so that the wire format can be the ADC's own 12-bit REAL samples at twice the IQ rate. ---- */
typedef struct sdrx_airspy_fe sdrx_airspy_fe;
#define SDRX_FE_PACKED12 0x1u   /* airspy_set_packing(1): 8 samples per three 32-bit words (airspy.c:323-338); else one
                                   sample per uint16 (0..4095)                                                     */
#define SDRX_FE_EXACT_DC 0x2u   /* run remove_dc (one recurrence over the whole stream) on a single thread: bit-exact by
                                   construction, ~4 ms per 640 000 samples; the default is segment-parallel          */
#define SDRX_FE_BESIDE_ENGINE 0x4u /* the front end feeds an sdrx_engine on the same GPU: remove_dc runs in its narrow form
                                   (512-sample segments, 5 CTAs, input staged through shared memory) so that it fits
                                   the SMs the channel bank's launch leaves free; same results, same 1e-6 bar       */
/* hb_kernel: HB_KERNEL_FLOAT (libairspy filters.h:28-30, 47 taps); max_samples: most REAL samples per call (multiple of 8) */
int   sdrx_airspy_fe_create(int device, const float *hb_kernel, int hb_len, unsigned max_samples, unsigned flags,
                            sdrx_airspy_fe **out);                                        /* iqconverter_float_create */
void  sdrx_airspy_fe_destroy(sdrx_airspy_fe *f);                                          /* iqconverter_float_free   */
int   sdrx_airspy_fe_reset(sdrx_airspy_fe *f);                                            /* iqconverter_float_reset  */
/* One libairspy buffer: n_samples REAL samples in (host pointer, wire format per flags; multiple of 8), n_samples/2
 * IQ pairs out.  Queued on the front end's stream; iq_host (may be NULL) receives the pairs and makes the call wait. */
int   sdrx_airspy_fe_process(sdrx_airspy_fe *f, const void *raw, unsigned n_samples, float *iq_host);
int   sdrx_airspy_fe_process_device(sdrx_airspy_fe *f, const void *d_raw, unsigned n_samples);
/* Device pointer to the last call's IQ pairs (float re,im): feed it to sdrx_engine_push_block_device(SDRX_FMT_F32)
 * after sdrx_engine_wait_event(e, sdrx_airspy_fe_done_event(f)) -- no host synchronisation in between.            */
const float *sdrx_airspy_fe_output(sdrx_airspy_fe *f, unsigned *n_pairs);
void *sdrx_airspy_fe_done_event(sdrx_airspy_fe *f);    /* cudaEvent_t */
void *sdrx_airspy_fe_stream(sdrx_airspy_fe *f);        /* cudaStream_t */
int   sdrx_airspy_fe_sync(sdrx_airspy_fe *f);
int   sdrx_airspy_fe_avg(sdrx_airspy_fe *f, float *avg);                                  /* cnv->avg */
uint64_t sdrx_airspy_fe_launch_count(sdrx_airspy_fe *f);
/* Orders the engine's ingest stream behind a CUDA event recorded by another producer (e.g. the front end above). */
int   sdrx_engine_wait_event(sdrx_engine *e, void *cuda_event);
/* ... and the other way round: the event after which the buffer given to the last push_block_device is free again;
 * the front end's output ring is two deep, so handing it to sdrx_airspy_fe_wait_event before every call keeps a
 * streaming producer safe: the NEXT call's output write (its last pass only) is ordered behind the event, which must
 * stay alive until that call has been made.                                                                        */
void *sdrx_engine_ingest_event(sdrx_engine *e);        /* cudaEvent_t, NULL before the first push */
int   sdrx_airspy_fe_wait_event(sdrx_airspy_fe *f, void *cuda_event);

#ifdef __cplusplus
}
#endif
#endif /* SDRX_B200_H */

/*
 * TEST INFRASTRUCTURE (oracle) -- not part of the product path.
 *
 * Plain-C CPU restatement of the sdrx per-channel DSP chain (reference: johanhedin/sdrx,
 * /root/reference/src).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (sdrx_b200/csrc) never does.
 *
 * Parity status: every function here is pinned BIT-EXACT against the reference's own
 * headers compiled in this container (oracle/_ref/libsdrx_ref.so, built by oracle/Makefile
 * from /root/reference/src) by tests/test_oracle.py, and against the golden vectors under
 * tests/golden/ that were generated from that library.  The one exception is the 512-point
 * squelch FFT: the reference calls FFTW3f, an un-vendored, un-pinned system library, so the
 * FFT boundary is "parity unpinned" (see fft512.h).
 *
 * Build: gcc -std=gnu11 -O2 -ffp-contract=off (no -march=native): every multiply and add
 * rounds separately, sums run strictly in the reference's portable order.
 *
 * All complex data are interleaved (re,im) float pairs == std::complex<float> == iqsample_t
 * (/root/reference/src/iqsample.hpp:26).
 */
#ifndef SDRX_ORACLE_H
#define SDRX_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- set-up helpers (sdrx.cpp) ---- */
/* sdrx.cpp:1682-1690 translator table; writes 2*N floats (nothing if ch_offset == 0). */
void sdrxo_translator(int N, int z, int ch_offset, float *out);
/* sdrx.cpp:898-900 Hamming window, 512 floats. */
void sdrxo_window(float *out);
/* sdrx.cpp:1024-1044 */
int  sdrxo_audio_pos(unsigned channel_no, unsigned num_channels);
/* sdrx.cpp:1449-1478: offset in 8.33 kHz steps of an aeronautical designator; -100000 on a bad name. */
int  sdrxo_channel_to_offset(const char *channel, int32_t tuner_fq);
/* sdrx.cpp:962-1019 */
uint32_t sdrxo_parse_fq(const char *str, int aeronautical);

/* ---- MSD (msd.hpp) ---- */
typedef struct sdrxo_msd sdrxo_msd;
sdrxo_msd *sdrxo_msd_create(const float *translator, int N, int n_stages, const int *m,
                            const int *ntaps, const float *taps, int use_ftfir);
sdrxo_msd *sdrxo_msd_clone(const sdrxo_msd *);
void   sdrxo_msd_destroy(sdrxo_msd *);
size_t sdrxo_msd_m(const sdrxo_msd *);
size_t sdrxo_msd_decimate(sdrxo_msd *, const float *in, size_t in_len, float *out);

/* ---- AGC / LfAGC (agc.hpp) ---- */
typedef struct {
    float attack, decay, reference, min_gain, max_gain, alpha, gain, power;
} sdrxo_agc;
typedef struct {
    float attack, decay, reference, min_gain, max_gain, alpha, gain, power;
    int   active;
} sdrxo_lfagc;
void  sdrxo_agc_init(sdrxo_agc *, float attack, float decay, float reference, float max_gain, float alpha);
void  sdrxo_agc_set_max_gain(sdrxo_agc *, float max_gain);
void  sdrxo_agc_adjust(sdrxo_agc *, const float *in, unsigned n, float *out);
void  sdrxo_lfagc_init(sdrxo_lfagc *, float attack, float decay, float reference, float max_gain, float alpha);
void  sdrxo_lfagc_set_max_gain(sdrxo_lfagc *, float max_gain);
void  sdrxo_lfagc_adjust(sdrxo_lfagc *, const float *in, unsigned n, float *out);

/* ---- FIR / FIR2 / FIR3<iqsample_t> (fir.hpp) ---- */
typedef struct sdrxo_fir sdrxo_fir;
/* kind: 1 = FIR (mono real), 2 = FIR2 (stereo interleaved), 3 = FIR3<iqsample_t> (complex) */
sdrxo_fir *sdrxo_fir_create(int kind, const float *c, int n);
void  sdrxo_fir_destroy(sdrxo_fir *);
void  sdrxo_fir_set_gain(sdrxo_fir *, float gain_db);
/* in_len follows the reference: floats for kind 1 and 2 (so 2*frames for FIR2), complex samples for 3 */
void  sdrxo_fir_filter(sdrxo_fir *, const float *in, unsigned in_len, float *out);

/* ---- whole receiver chain ---- */
typedef struct sdrxo_rx sdrxo_rx;
typedef struct {
    int          n_channels;
    int          N;                 /* translator length                                 */
    const float *translators;       /* [C][N][2]; a channel whose table is all-zero length
                                       is flagged by has_translator[c] == 0              */
    const uint8_t *has_translator;  /* [C]                                               */
    int          n_stages;
    const int   *m;                 /* [n_stages] */
    const int   *ntaps;             /* [n_stages] */
    const float *taps;              /* concatenated */
    const float *sql_level;         /* [C] */
    const int   *pos;               /* [C] audio position -2..2 */
    int          use_lf_agc;
    int          use_ftfir;
    float        lf_gain_db;
    const float *fir2_taps; int fir2_ntaps;   /* coeff_bp4am_channel */
    const float *passband_shape;    /* [512] */
    const float *ramp_up;           /* [512] */
    const float *ramp_down;         /* [512] */
} sdrxo_rx_config;

sdrxo_rx *sdrxo_rx_create(const sdrxo_rx_config *cfg);
void  sdrxo_rx_destroy(sdrxo_rx *);
/* fmt 0 u8 (rtl_dev.cpp:316-327), 1 int16 (float = s16/32768), 2 float32; returns pwr_dbfs
 * (rtl_dev.cpp:329-341, airspy_dev.cpp:394-405). */
float sdrxo_rx_ingest(sdrxo_rx *, const void *raw, int fmt, unsigned n_samples);
/* sdrx.cpp:342-345 */
void  sdrxo_rx_channelize(sdrxo_rx *, float *ch_iq_out);
/* sdrx.cpp:452-573, :646-656 */
void  sdrxo_rx_demod(sdrxo_rx *, float *audio, uint8_t *sql_state, float *snr,
                     float *mix_pre, float *mix, int16_t *pcm);
void  sdrxo_rx_gains(sdrxo_rx *, float *agc_gain, float *lf_gain);
/* per-channel modulation, 0 AM (default) / 1 FM: Channel::mod, sdrx.cpp:230, :1351-1358 */
void  sdrxo_rx_set_modulation(sdrxo_rx *, const int *mod);

#ifdef __cplusplus
}
#endif
#endif

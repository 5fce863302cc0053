// TEST INFRASTRUCTURE (oracle/_ref) -- not part of the product path.
//
// Thin C-ABI harness around the REFERENCE's own headers, compiled from where they lie
// under /root/reference/src (see oracle/Makefile; nothing is copied into this repo):
//   iqsample.hpp, msd.hpp, ds.hpp, fir.hpp, agc.hpp, coeffs.hpp, filters/*.hpp
// The classes MSD, DS, AGC, LfAGC, FIR, FIR2, FIR3 and all coefficient tables used below
// are the reference's, unmodified.  Only the glue of sdrx.cpp -- which cannot be built
// here (popt, fftw3f, alsa, sigc++, libusb are absent) -- is restated, statement for
// statement, each block citing the lines it follows.  FFTW3f is replaced by
// oracle/fft512.c (parity unpinned at that boundary, see fft512.h).
//
// Built twice: portable (-O2 -ffp-contract=off: defines parity) and with the
// reference's Release flags (-O3 -DNDEBUG + AVX2/FMA: the timed CPU baseline).

#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>
#include <memory>
#include <latch>
#include <thread>
#include <complex>

#include <iqsample.hpp>
#include <msd.hpp>
#include <ds.hpp>
#include <fir.hpp>
#include <agc.hpp>
#include <coeffs.hpp>
#include <filters/fs_00960_08bit_ds_to_00016.hpp>
#include <filters/fs_01200_08bit_ds_to_00016.hpp>
#include <filters/fs_01440_08bit_ds_to_00016.hpp>
#include <filters/fs_01600_08bit_ds_to_00016.hpp>
#include <filters/fs_01920_08bit_ds_to_00016.hpp>
#include <filters/fs_02400_08bit_ds_to_00016.hpp>
#include <filters/fs_02560_08bit_ds_to_00016.hpp>
#include <filters/fs_06000_12bit_ds_to_00016.hpp>
#include <filters/fs_10000_12bit_ds_to_00016.hpp>
#include <filters/fs_00016_16bit_ch.hpp>

#include "fft512.h"

#define CH_IQ_BUF_SIZE 512   // sdrx.cpp:129
#define FFT_SIZE CH_IQ_BUF_SIZE

namespace {

// sdrx.cpp:1584-1677 -- N, z and the stage list per sample rate.
bool rate_config(uint32_t rate, int &N, int &z, std::vector<MSD::Stage> &stages) {
    switch (rate) {
        case 960000:
            N = 576; z = 5;
            stages = { {3, fs_00960_08bit_ds_lpf1_00960_to_00320},
                       {4, fs_00960_08bit_ds_lpf2_00320_to_00080},
                       {5, fs_00960_08bit_ds_lpf3_00080_to_00016} };
            break;
        case 1200000:
            N = 144; z = 1;
            stages = { {3, fs_01200_08bit_ds_lpf1_01200_to_00400},
                       {5, fs_01200_08bit_ds_lpf2_00400_to_00080},
                       {5, fs_01200_08bit_ds_lpf3_00080_to_00016} };
            break;
        case 1440000:
            N = 1728; z = 10;
            stages = { {3, fs_01440_08bit_ds_lpf1_01440_to_00400},
                       {6, fs_01440_08bit_ds_lpf2_00480_to_00080},
                       {5, fs_01440_08bit_ds_lpf3_00080_to_00016} };
            break;
        case 1600000:
            N = 192; z = 1;
            stages = { {4, fs_01600_08bit_ds_lpf1_01600_to_00400},
                       {5, fs_01600_08bit_ds_lpf2_00400_to_00080},
                       {5, fs_01600_08bit_ds_lpf3_00080_to_00016} };
            break;
        case 1920000:
            N = 1152; z = 5;
            stages = { {4, fs_01920_08bit_ds_lpf1_01920_to_00480},
                       {6, fs_01920_08bit_ds_lpf2_00480_to_00080},
                       {5, fs_01920_08bit_ds_lpf3_00080_to_00016} };
            break;
        case 2400000:
            N = 288; z = 1;
            stages = { {2, fs_02400_08bit_ds_lpf1_02400_to_01200},
                       {3, fs_02400_08bit_ds_lpf2_01200_to_00400},
                       {5, fs_02400_08bit_ds_lpf3_00400_to_00080},
                       {5, fs_02400_08bit_ds_lpf4_00080_to_00016} };
            break;
        case 2560000:
            N = 1536; z = 5;
            stages = { {20, fs_02560_08bit_ds_lpf1_02560_to_00128},
                       { 4, fs_02560_08bit_ds_lpf2_00128_to_00032},
                       { 2, fs_02560_08bit_ds_lpf4_00032_to_00016} };
            break;
        case 6000000:
            N = 720; z = 1;
            stages = { {15, fs_06000_12bit_ds_lpf1_06000_to_00400},
                       { 5, fs_06000_12bit_ds_lpf3_00400_to_00080},
                       { 5, fs_06000_12bit_ds_lpf4_00080_to_00016} };
            break;
        case 10000000:
            N = 1200; z = 1;
            stages = { {5, fs_10000_12bit_ds_lpf1_10000_to_02000},
                       {5, fs_10000_12bit_ds_lpf2_02000_to_00400},
                       {5, fs_10000_12bit_ds_lpf3_00400_to_00080},
                       {5, fs_10000_12bit_ds_lpf4_00080_to_00016} };
            break;
        default:
            N = 0; z = 1; stages.clear();
            return false;
    }
    return true;
}

// sdrx.cpp:1682-1690 -- the translator ("tuner") table of one channel.
std::vector<iqsample_t> make_translator(int N, int z, int ch_offset) {
    std::vector<iqsample_t> translator;
    if (ch_offset != 0) {
        for (int n = 0; n < N; n++) {
            std::complex<float> e(0.0f, -2.0f * M_PI * n * ch_offset * (float)z/(float)N);
            translator.push_back(exp(e));
        }
    }
    return translator;
}

enum { SQL_CLOSED = 0, SQL_OPEN = 1 };   // sdrx.cpp:149

// The DSP members of sdrx.cpp:206-236 `class Channel`.
struct RefChannel {
    MSD    msd;
    std::unique_ptr<DS> ds;
    AGC    agc;
    LfAGC  agc_lf;
    float  sql_level = 9.0f;
    int    sql_state = SQL_CLOSED;
    int    sql_state_prev = SQL_CLOSED;
    int    pos = 0;
    int    mod = 0;                       // 0 AM, 1 FM (Channel::mod, sdrx.cpp:230)
    iqsample_t prev{0.0f, 0.0f};          // Demod::prev_sample (sdrx.cpp:202)
};

// Demod::demod (sdrx.cpp:178-200) lives inside sdrx.cpp next to main() and cannot be included; this is the same
// arithmetic on the reference's own iqsample_t (std::complex<float>) so that the C restatement's reading of
// "complex / float" and of the atan2 arguments is checked against what the C++ library does.
static float ref_demod(RefChannel &ch, iqsample_t v) {
    if (ch.mod != 1) return std::abs(v);
    const iqsample_t u = v / std::abs(v);
    const float cross = u.imag() * ch.prev.real() - u.real() * ch.prev.imag();
    const float dot = u.real() * ch.prev.real() + u.imag() * ch.prev.imag();
    ch.prev = u;
    return std::atan2(cross, dot);
}

struct RefRx {
    uint32_t rate = 0;
    int      N = 0, z = 1;
    std::vector<MSD::Stage> stages;
    std::vector<RefChannel> channels;
    FIR2     audio_filter;
    float    window[FFT_SIZE];
    iqsample_t fft_in[FFT_SIZE];
    iqsample_t fft_out[FFT_SIZE];
    std::vector<iqsample_t> iq_in;      // device-layer float buffer (rtl_dev.hpp:74 iq_buffer_)
    std::vector<iqsample_t> chunk;      // one CRB chunk [C][512]
    bool     threaded = false;
};

// sdrx.cpp:1024-1044
int get_audio_pos(unsigned channel_no, unsigned num_channels) {
    unsigned num_positions = 5;
    unsigned half = num_channels / 2;
    unsigned odd = (num_channels % 2) == 0 ? 0 : 1;
    float    tmp;
    int      pos = 0;

    if (channel_no < num_channels) {
        if (channel_no < half) {
            tmp = (float)(channel_no * num_positions) / (float)num_channels;
            pos = floorf(tmp) - num_positions/2;
        } else if (channel_no == half && odd) {
            pos = 0;
        } else {
            tmp = (float)((num_channels-1 - channel_no)*num_positions) / (float)num_channels;
            pos = num_positions/2 - floorf(tmp);
        }
    }
    return pos;
}

inline void pan_mix(float *mix, unsigned i, int pos, float s) {
    // sdrx.cpp:471-497 (identical again at :505-531)
    switch (pos) {
        case -2: mix[i*2] += 0.8f*s; mix[i*2+1] += 0.2f*s; break;
        case -1: mix[i*2] += 0.6f*s; mix[i*2+1] += 0.4f*s; break;
        case  1: mix[i*2] += 0.4f*s; mix[i*2+1] += 0.6f*s; break;
        case  2: mix[i*2] += 0.2f*s; mix[i*2+1] += 0.8f*s; break;
        default: mix[i*2] += 0.5f*s; mix[i*2+1] += 0.5f*s; break;
    }
}

} // namespace

extern "C" {

// ---------------------------------------------------------------- tables
int sdrxref_rate_config(uint32_t rate, int *N, int *z, int *n_stages, int *m, int *ntaps,
                        float *taps, int taps_cap) {
    int n, zz; std::vector<MSD::Stage> st;
    if (!rate_config(rate, n, zz, st)) return -1;
    *N = n; *z = zz; *n_stages = (int)st.size();
    int pos = 0;
    for (size_t s = 0; s < st.size(); ++s) {
        m[s] = (int)st[s].m; ntaps[s] = (int)st[s].h.size();
        if (pos + ntaps[s] > taps_cap) return -2;
        memcpy(taps + pos, st[s].h.data(), sizeof(float) * st[s].h.size());
        pos += ntaps[s];
    }
    return 0;
}

int sdrxref_table(const char *name, float *out, int cap) {
    const std::vector<float> *v = nullptr;
    std::string n(name);
    if      (n == "coeff_bp4am_channel")            v = &coeff_bp4am_channel;
    else if (n == "passband_shape")                 v = &passband_shape;
    else if (n == "ramp_up")                        v = &ramp_up;
    else if (n == "ramp_down")                      v = &ramp_down;
    else if (n == "fs_00016_16bit_ch_amdemod_lpf1") v = &fs_00016_16bit_ch_amdemod_lpf1;
    if (!v) return -1;
    if (out) { if ((int)v->size() > cap) return -2; memcpy(out, v->data(), sizeof(float) * v->size()); }
    return (int)v->size();
}

void sdrxref_translator(int N, int z, int ch_offset, float *out) {
    auto t = make_translator(N, z, ch_offset);
    if (!t.empty()) memcpy(out, t.data(), sizeof(iqsample_t) * t.size());
}

// sdrx.cpp:898-900
void sdrxref_window(float *out) {
    for (unsigned n = 0; n < FFT_SIZE; n++) {
        out[n] = 0.54f - 0.46f * std::cos((2.0f * M_PI * n) / FFT_SIZE);
    }
}

int sdrxref_audio_pos(unsigned ch, unsigned n) { return get_audio_pos(ch, n); }

// ---------------------------------------------------------------- MSD (msd.hpp)
void *sdrxref_msd_create(const float *translator, int N, int n_stages, const int *m,
                         const int *ntaps, const float *taps, int use_ftfir) {
    std::vector<iqsample_t> tr;
    for (int i = 0; i < N; ++i) tr.emplace_back(translator[2*i], translator[2*i+1]);
    std::vector<MSD::Stage> st;
    int pos = 0;
    for (int s = 0; s < n_stages; ++s) {
        st.push_back(MSD::Stage{ (size_t)m[s], std::vector<float>(taps + pos, taps + pos + ntaps[s]) });
        pos += ntaps[s];
    }
    return new MSD(tr, st, use_ftfir != 0);
}
void *sdrxref_msd_clone(void *p) { return new MSD(*reinterpret_cast<MSD*>(p)); }
void sdrxref_msd_destroy(void *p) { delete reinterpret_cast<MSD*>(p); }
size_t sdrxref_msd_m(void *p) { return reinterpret_cast<MSD*>(p)->m(); }
size_t sdrxref_msd_decimate(void *p, const float *in, size_t in_len, float *out) {
    size_t out_len = 0;
    reinterpret_cast<MSD*>(p)->decimate(reinterpret_cast<const iqsample_t*>(in), in_len,
                                        reinterpret_cast<iqsample_t*>(out), &out_len);
    return out_len;
}

// DS (ds.hpp): run `n` MSD copies as threaded downsamplers over one shared block, the
// way sdrx.cpp:327-340 drives them.  msds[i] is cloned into the DS (ds.hpp:34), so the
// caller's MSD state is NOT advanced -- exactly the reference's semantics.
struct RefDsBank { std::vector<std::unique_ptr<DS>> ds; };
void *sdrxref_dsbank_create(void **msds, int n) {
    auto *b = new RefDsBank;
    for (int i = 0; i < n; ++i) b->ds.emplace_back(new DS(*reinterpret_cast<MSD*>(msds[i])));
    return b;
}
void sdrxref_dsbank_destroy(void *p) { delete reinterpret_cast<RefDsBank*>(p); }
void sdrxref_dsbank_run(void *p, const float *in, unsigned in_len, float *out /*[n][512]*/) {
    auto *b = reinterpret_cast<RefDsBank*>(p);
    auto latch = std::make_shared<std::latch>(b->ds.size());
    iqsample_t *iq_buf_ptr = reinterpret_cast<iqsample_t*>(out);
    for (auto &d : b->ds) {
        d->addJob(reinterpret_cast<const iqsample_t*>(in), in_len, iq_buf_ptr, latch);
        iq_buf_ptr += CH_IQ_BUF_SIZE;
    }
    latch->wait();
}

// ---------------------------------------------------------------- AGC / LfAGC (agc.hpp)
void *sdrxref_agc_create(float attack, float decay, float reference, float max_gain, float alpha) {
    return new AGC(attack, decay, reference, max_gain, alpha);
}
void sdrxref_agc_destroy(void *p) { delete reinterpret_cast<AGC*>(p); }
void sdrxref_agc_set(void *p, float attack, float decay, float reference, float max_gain) {
    auto *a = reinterpret_cast<AGC*>(p);
    a->setReference(reference); a->setAttack(attack); a->setDecay(decay); a->setMaxGain(max_gain);
}
float sdrxref_agc_gain(void *p) { return reinterpret_cast<AGC*>(p)->gain(); }
void sdrxref_agc_adjust(void *p, const float *in, unsigned n, float *out) {
    auto *a = reinterpret_cast<AGC*>(p);
    for (unsigned i = 0; i < n; ++i) {
        iqsample_t y = a->adjust(iqsample_t(in[2*i], in[2*i+1]));
        out[2*i] = y.real(); out[2*i+1] = y.imag();
    }
}
void *sdrxref_lfagc_create(float attack, float decay, float reference, float max_gain, float alpha) {
    return new LfAGC(attack, decay, reference, max_gain, alpha);
}
void sdrxref_lfagc_destroy(void *p) { delete reinterpret_cast<LfAGC*>(p); }
void sdrxref_lfagc_set(void *p, float attack, float decay, float reference, int active) {
    auto *a = reinterpret_cast<LfAGC*>(p);
    a->setReference(reference); a->setAttack(attack); a->setDecay(decay);
    if (active) a->activate(); else a->deactivate();
}
float sdrxref_lfagc_gain(void *p) { return reinterpret_cast<LfAGC*>(p)->gain(); }
void sdrxref_lfagc_adjust(void *p, const float *in, unsigned n, float *out) {
    auto *a = reinterpret_cast<LfAGC*>(p);
    for (unsigned i = 0; i < n; ++i) out[i] = a->adjust(in[i]);
}

// ---------------------------------------------------------------- FIR / FIR2 / FIR3 (fir.hpp)
void *sdrxref_fir_create(const float *c, int n) { return new FIR(std::vector<float>(c, c + n)); }
void sdrxref_fir_destroy(void *p) { delete reinterpret_cast<FIR*>(p); }
void sdrxref_fir_set_gain(void *p, float db) { reinterpret_cast<FIR*>(p)->setGain(db); }
void sdrxref_fir_filter(void *p, const float *in, unsigned n, float *out) { reinterpret_cast<FIR*>(p)->filter(in, n, out); }

void *sdrxref_fir2_create(const float *c, int n) { return new FIR2(std::vector<float>(c, c + n)); }
void sdrxref_fir2_destroy(void *p) { delete reinterpret_cast<FIR2*>(p); }
void sdrxref_fir2_set_gain(void *p, float db) { reinterpret_cast<FIR2*>(p)->setGain(db); }
void sdrxref_fir2_filter(void *p, const float *in, unsigned n, float *out) { reinterpret_cast<FIR2*>(p)->filter(in, n, out); }

void *sdrxref_fir3c_create(const float *c, int n) { return new FIR3<iqsample_t>(std::vector<float>(c, c + n)); }
void sdrxref_fir3c_destroy(void *p) { delete reinterpret_cast<FIR3<iqsample_t>*>(p); }
void sdrxref_fir3c_set_gain(void *p, float db) { reinterpret_cast<FIR3<iqsample_t>*>(p)->setGain(db); }
void sdrxref_fir3c_filter(void *p, const float *in, unsigned n, float *out) {
    reinterpret_cast<FIR3<iqsample_t>*>(p)->filter(reinterpret_cast<const iqsample_t*>(in), n,
                                                   reinterpret_cast<iqsample_t*>(out));
}

// ---------------------------------------------------------------- whole receiver chain
// Set-up: sdrx.cpp:1584-1714 (rate tables, per-channel translator/MSD/AGC/LfAGC/pos)
// and :880-881, :898-900 (FIR2, window).
void *sdrxref_rx_create(uint32_t rate, int n_channels, const int *ch_offsets, const float *sql_levels,
                        int use_lf_agc, float lf_gain_db, int use_ftfir, int threaded) {
    auto *rx = new RefRx;
    rx->rate = rate;
    if (!rate_config(rate, rx->N, rx->z, rx->stages)) { delete rx; return nullptr; }
    rx->channels.resize(n_channels);
    unsigned ch_idx = 0;
    for (auto &ch : rx->channels) {
        std::vector<iqsample_t> translator = make_translator(rx->N, rx->z, ch_offsets[ch_idx]);
        ch.msd = MSD(translator, rx->stages, use_ftfir != 0);
        if (threaded) ch.ds.reset(new DS(ch.msd));

        ch.agc.setReference(1.0f);
        ch.agc.setAttack(1.0f);
        ch.agc.setDecay(0.01f);
        ch.agc.setMaxGain(300);

        ch.agc_lf.setReference(1.0f);
        ch.agc_lf.setAttack(1.0f);
        ch.agc_lf.setDecay(0.01f);
        if (use_lf_agc) ch.agc_lf.activate();

        ch.sql_level = sql_levels ? sql_levels[ch_idx] : 9.0f;
        ch.pos = get_audio_pos(ch_idx, rx->channels.size());
        ++ch_idx;
    }
    rx->audio_filter = FIR2(coeff_bp4am_channel);
    rx->audio_filter.setGain(lf_gain_db);
    sdrxref_window(rx->window);
    rx->chunk.assign((size_t)n_channels * CH_IQ_BUF_SIZE, iqsample_t(0.0f, 0.0f));
    rx->threaded = threaded != 0;
    return rx;
}
void sdrxref_rx_destroy(void *p) { delete reinterpret_cast<RefRx*>(p); }
int sdrxref_rx_pos(void *p, int ch) { return reinterpret_cast<RefRx*>(p)->channels[ch].pos; }

// Device layer: raw -> iqsample_t and block power.
//   fmt 0: RTL u8        rtl_dev.cpp:316-341
//   fmt 1: int16 IQ      libairspy convention float = s16/32768 (airspy.c:58-62, :304)
//   fmt 2: float32 IQ    airspy_dev.cpp:355-357 (copy), power :394-405 (portable variant)
float sdrxref_ingest(void *p, const void *raw, int fmt, unsigned n_samples) {
    auto *rx = reinterpret_cast<RefRx*>(p);
    rx->iq_in.resize(n_samples);
    iqsample_t *iq = rx->iq_in.data();
    if (fmt == 0) {
        const unsigned char *data = reinterpret_cast<const unsigned char*>(raw);
        uint32_t data_pos = 0; unsigned iq_pos = 0;
        while (iq_pos < n_samples) {
            iq[iq_pos].real(((float)data[data_pos])   / 127.5f - 1.0f);
            iq[iq_pos].imag(((float)data[data_pos+1]) / 127.5f - 1.0f);
            iq_pos++; data_pos += 2;
        }
    } else if (fmt == 1) {
        const int16_t *data = reinterpret_cast<const int16_t*>(raw);
        for (unsigned i = 0; i < n_samples; ++i)
            iq[i] = iqsample_t((float)data[2*i] * (1.0f/32768.0f), (float)data[2*i+1] * (1.0f/32768.0f));
    } else {
        const float *data = reinterpret_cast<const float*>(raw);
        int sample_pos = 0;
        for (unsigned i = 0; i < n_samples; ++i) {
            iq[i] = iqsample_t(data[sample_pos], data[sample_pos+1]);
            sample_pos += 2;
        }
    }
    float pwr_rms = 0.0f;
    for (unsigned i = 0; i < n_samples; i++) {
        float ampl_squared = std::norm(iq[i]);
        pwr_rms += ampl_squared;
    }
    pwr_rms = pwr_rms / n_samples;
    return 10 * std::log10(pwr_rms) - 3.0f;
}

void sdrxref_rx_set_modulation(void *p, const int *mod) {
    auto *rx = reinterpret_cast<RefRx*>(p);
    unsigned c = 0;
    for (auto &ch : rx->channels) { ch.mod = mod ? mod[c] : 0; ++c; }
}

// Input side, sdrx.cpp:327-346: channelize the block already converted by
// sdrxref_ingest() into the chunk [C][512].
void sdrxref_rx_channelize(void *p, float *ch_iq_out /* [C][512][2] or NULL */) {
    auto *rx = reinterpret_cast<RefRx*>(p);
    const iqsample_t *data = rx->iq_in.data();
    unsigned data_len = (unsigned)rx->iq_in.size();
    iqsample_t *iq_buf_ptr = rx->chunk.data();
    if (rx->threaded) {
        auto latch = std::make_shared<std::latch>(rx->channels.size());
        for (auto &ch : rx->channels) {
            ch.ds->addJob(data, data_len, iq_buf_ptr, latch);
            iq_buf_ptr += CH_IQ_BUF_SIZE;
        }
        latch->wait();
    } else {
        for (auto &ch : rx->channels) {
            ch.msd.decimate(data, data_len, iq_buf_ptr);
            iq_buf_ptr += CH_IQ_BUF_SIZE;
        }
    }
    if (ch_iq_out) memcpy(ch_iq_out, rx->chunk.data(), sizeof(iqsample_t) * rx->chunk.size());
}

// Output side, sdrx.cpp:452-573 and :646-656, on the chunk produced by
// sdrxref_rx_channelize().  Outputs (any may be NULL):
//   audio[C][512]   the per-channel sample `s` that is mixed (0 where nothing is mixed)
//   sql_state[C]    state AFTER this block's decision;  snr[C]
//   mix_pre[1024]   stereo mix before FIR2; mix[1024] after FIR2; pcm[1024] int16
void sdrxref_rx_demod(void *p, float *audio, uint8_t *sql_state, float *snr_out,
                      float *mix_pre, float *mix, int16_t *pcm) {
    auto *rx = reinterpret_cast<RefRx*>(p);
    const iqsample_t *iq_buffer = rx->chunk.data();
    float   audio_buffer_float[CH_IQ_BUF_SIZE*2];
    int16_t audio_buffer_s16[CH_IQ_BUF_SIZE*2];

    memset(audio_buffer_float, 0, sizeof(audio_buffer_float));
    if (audio) memset(audio, 0, sizeof(float) * rx->channels.size() * CH_IQ_BUF_SIZE);
    unsigned j = 0;
    unsigned c = 0;
    for (auto &ch : rx->channels) {
        for (unsigned i = 0; i < CH_IQ_BUF_SIZE; ++i) {
            iqsample_t agc_adj_sample = ch.agc.adjust(iq_buffer[j]);
            if (ch.sql_state == SQL_OPEN) {
                float s = ref_demod(ch, agc_adj_sample);     // Demod::demod, sdrx.cpp:178-200
                s = ch.agc_lf.adjust(s);
                if (ch.sql_state_prev == SQL_CLOSED) {
                    s = ramp_up[i] * s;
                }
                pan_mix(audio_buffer_float, i, ch.pos, s);
                if (audio) audio[(size_t)c * CH_IQ_BUF_SIZE + i] = s;
            } else if (ch.sql_state_prev == SQL_OPEN) {
                float s = std::abs(agc_adj_sample);
                s = ch.agc_lf.adjust(s);
                s = ramp_down[i] * s;
                pan_mix(audio_buffer_float, i, ch.pos, s);
                if (audio) audio[(size_t)c * CH_IQ_BUF_SIZE + i] = s;
            }
            rx->fft_in[i] = iq_buffer[j] * rx->window[i];
            ++j;
        }
        ch.sql_state_prev = ch.sql_state;

        sdrx_oracle_fft512(reinterpret_cast<const float*>(rx->fft_in), reinterpret_cast<float*>(rx->fft_out));

        float sig_level = 0.0f;
        for (unsigned i = 3; i < 91; i++) {
            sig_level += std::norm(rx->fft_out[i]);
            sig_level += std::norm(rx->fft_out[FFT_SIZE - i]);
        }
        sig_level /= 176;

        float ref_level_hi = 0.0f;
        float ref_level_lo = 0.0f;
        for (unsigned i = 112; i < 157; i++) {
            ref_level_hi += std::norm(rx->fft_out[i] * passband_shape[i]);
            ref_level_lo += std::norm(rx->fft_out[FFT_SIZE - i] * passband_shape[FFT_SIZE - i]);
        }
        ref_level_hi /= 45;
        ref_level_lo /= 45;
        float noise_level = (ref_level_hi + ref_level_lo) / 2;

        float snr = 10 * std::log10(sig_level / noise_level);

        if (snr > ch.sql_level + 3 || ch.sql_level == 0.0f) {
            ch.sql_state = SQL_OPEN;
        } else if (snr < ch.sql_level) {
            ch.sql_state = SQL_CLOSED;
        }
        if (sql_state) sql_state[c] = (uint8_t)ch.sql_state;
        if (snr_out) snr_out[c] = snr;
        ++c;
    }

    if (mix_pre) memcpy(mix_pre, audio_buffer_float, sizeof(audio_buffer_float));

    rx->audio_filter.filter(audio_buffer_float, CH_IQ_BUF_SIZE*2, audio_buffer_float);

    for (unsigned i = 0; i < CH_IQ_BUF_SIZE*2; ++i) {
        int16_t s;
        if (audio_buffer_float[i] > 1.0f)       s = 32767;
        else if (audio_buffer_float[i] < -1.0f) s = -32767;
        else s = (int16_t)(audio_buffer_float[i] * 32767.0f);
        audio_buffer_s16[i] = s;
    }
    if (mix) memcpy(mix, audio_buffer_float, sizeof(audio_buffer_float));
    if (pcm) memcpy(pcm, audio_buffer_s16, sizeof(audio_buffer_s16));
}

// AGC state peek (for state save/load parity)
void sdrxref_rx_gains(void *p, float *agc_gain, float *lf_gain) {
    auto *rx = reinterpret_cast<RefRx*>(p);
    for (size_t c = 0; c < rx->channels.size(); ++c) {
        if (agc_gain) agc_gain[c] = rx->channels[c].agc.gain();
        if (lf_gain)  lf_gain[c]  = rx->channels[c].agc_lf.gain();
    }
}

const char *sdrxref_build_info(void) {
#if defined(__AVX2__)
    return "reference headers, AVX2 path (msd.hpp:274-307)";
#else
    return "reference headers, portable path (msd.hpp:308-332)";
#endif
}

} // extern "C"

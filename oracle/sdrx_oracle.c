/*
 * TEST INFRASTRUCTURE (oracle) -- see sdrx_oracle.h for status and build flags.
 * Every function cites the reference lines (under /root/reference/src) it restates.
 */
#include "sdrx_oracle.h"
#include "fft512.h"

#include <complex.h>
#include <ctype.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define CH_IQ_BUF_SIZE 512 /* sdrx.cpp:129 */
#define FFT_SIZE 512       /* sdrx.cpp:130 */

typedef struct { float re, im; } cf;

/* ------------------------------------------------------------------ set-up helpers */

/* sdrx.cpp:1685-1690.  The angle expression is evaluated in double (M_PI is a double),
 * narrowed to float by the std::complex<float> constructor, then exp(complex<float>)
 * == glibc cexpf. */
void sdrxo_translator(int N, int z, int ch_offset, float *out) {
    if (ch_offset == 0) return;
    for (int n = 0; n < N; n++) {
        float a = (float)(-2.0f * M_PI * n * ch_offset * (float)z / (float)N);
        float complex e = cexpf(CMPLXF(0.0f, a));
        out[2 * n + 0] = crealf(e);
        out[2 * n + 1] = cimagf(e);
    }
}

/* sdrx.cpp:898-900 */
void sdrxo_window(float *out) {
    for (unsigned n = 0; n < FFT_SIZE; n++)
        out[n] = (float)(0.54f - 0.46f * cos((2.0f * M_PI * n) / FFT_SIZE));
}

/* sdrx.cpp:1024-1044 */
int sdrxo_audio_pos(unsigned channel_no, unsigned num_channels) {
    unsigned num_positions = 5;
    unsigned half = num_channels / 2;
    unsigned odd = (num_channels % 2) == 0 ? 0 : 1;
    float tmp;
    int pos = 0;
    if (channel_no < num_channels) {
        if (channel_no < half) {
            tmp = (float)(channel_no * num_positions) / (float)num_channels;
            pos = (int)(floorf(tmp) - num_positions / 2);
        } else if (channel_no == half && odd) {
            pos = 0;
        } else {
            tmp = (float)((num_channels - 1 - channel_no) * num_positions) / (float)num_channels;
            pos = (int)(num_positions / 2 - floorf(tmp));
        }
    }
    return pos;
}

/* The two-digit sub-channel maps of sdrx.cpp:991-996 (Hz) and :1464-1469 (8.33 kHz steps). */
static const struct { const char *d; unsigned hz; int step; } sub_ch_map[] = {
    {"00", 0, 0},      {"05", 0, 0},      {"10", 8333, 1},   {"15", 16667, 2},
    {"25", 25000, 3},  {"30", 25000, 3},  {"35", 33333, 4},  {"40", 41667, 5},
    {"50", 50000, 6},  {"55", 50000, 6},  {"60", 58333, 7},  {"65", 66667, 8},
    {"75", 75000, 9},  {"80", 75000, 9},  {"85", 83333, 10}, {"90", 91667, 11},
};

static int sub_lookup(const char *two) {
    for (size_t i = 0; i < sizeof(sub_ch_map) / sizeof(sub_ch_map[0]); ++i)
        if (two[0] == sub_ch_map[i].d[0] && two[1] == sub_ch_map[i].d[1]) return (int)i;
    return -1;
}

static int all_digits(const char *s, size_t n) {
    for (size_t i = 0; i < n; ++i)
        if (!isdigit((unsigned char)s[i])) return 0;
    return 1;
}

/* sdrx.cpp:962-1019 */
uint32_t sdrxo_parse_fq(const char *str, int aeronautical) {
    uint32_t fq = 0;
    unsigned mhz = 0, hz = 0;
    const char *dot = strchr(str, '.');
    if (!dot) return 0;
    size_t int_len = (size_t)(dot - str);
    const char *frac = dot + 1;
    size_t frac_len = strlen(frac);
    if (!all_digits(str, int_len) || !all_digits(frac, frac_len) || int_len < 2 || int_len > 4 ||
        frac_len == 0 || frac_len > 6)
        return 0;
    if (aeronautical && frac_len != 3) return 0;
    char int_str[8];
    memcpy(int_str, str, int_len);
    int_str[int_len] = 0;
    if (aeronautical) {
        int k = sub_lookup(frac + 1);
        if (k >= 0) {
            mhz = (unsigned)atol(int_str);
            hz = (unsigned)(frac[0] - '0') * 100000 + sub_ch_map[k].hz;
        }
    } else {
        static const unsigned mult[6] = {100000, 10000, 1000, 100, 10, 1};
        mhz = (unsigned)atol(int_str);
        for (size_t i = 0; i < frac_len; ++i) hz += (unsigned)(frac[i] - '0') * mult[i];
    }
    if (mhz < 4000) fq = mhz * 1000000 + hz;
    return fq;
}

/* sdrx.cpp:1449-1478 */
int sdrxo_channel_to_offset(const char *channel, int32_t tuner_fq) {
    const char *dot = strchr(channel, '.');
    if (!dot || strlen(dot + 1) != 3) return -100000;
    char int_str[16];
    size_t int_len = (size_t)(dot - channel);
    if (int_len >= sizeof(int_str)) return -100000;
    memcpy(int_str, channel, int_len);
    int_str[int_len] = 0;
    const char *frac = dot + 1;
    int k = sub_lookup(frac + 1);
    if (k < 0) return -100000;
    int32_t fq_base = (int32_t)atol(int_str) * 1000000;
    fq_base += (frac[0] - '0') * 100000;
    int32_t fq_diff = fq_base - tuner_fq;
    int32_t offset_diff = (fq_diff / 100000) * 12;
    return offset_diff + sub_ch_map[k].step;
}

/* ------------------------------------------------------------------ MSD (msd.hpp) */

typedef struct {
    size_t m;      /* msd.hpp:428 m_   */
    size_t H;      /* h_.size()        */
    float *h;      /* msd.hpp:429 h_ (real value; the reference stores (h,h) pairs, :182-187) */
    cf    *d;      /* msd.hpp:430 d_, 2*H */
    size_t pos;    /* msd.hpp:431 */
    size_t isn;    /* msd.hpp:432 */
    cf    *hk;     /* msd.hpp:433 hk_ [num_sets][H] */
    size_t num_sets;
    size_t k;      /* msd.hpp:434 */
} stage_t;

struct sdrxo_msd {
    int      n_stages;
    stage_t *st;
    size_t   m_total;
    cf      *translator;
    size_t   N;
    size_t   trans_pos;
    int      use_ftfir;
};

/* msd.hpp:166-225 (S::S) */
static void stage_init(stage_t *s, size_t m, const float *h, size_t H, const cf *translator, size_t N) {
    s->m = m; s->H = H;
    s->h = (float *)malloc(sizeof(float) * H);
    memcpy(s->h, h, sizeof(float) * H);
    s->d = (cf *)calloc(2 * H, sizeof(cf));
    s->pos = 0; s->isn = m; s->k = 0;
    s->hk = NULL; s->num_sets = 0;
    if (translator && N > 0) {
        size_t num_sets = N / m;
        s->num_sets = num_sets;
        s->hk = (cf *)malloc(sizeof(cf) * num_sets * H);
        size_t j = (N - (H - 1) % N) % N;
        for (size_t set = 0; set < num_sets; set++) {
            size_t k = j;
            for (size_t i = 0; i < H; ++i) {
                /* *iter * translator[k] * 2.0f : float*complex, then complex*float */
                cf c;
                c.re = translator[k].re * h[i];
                c.im = translator[k].im * h[i];
                c.re = c.re * 2.0f;
                c.im = c.im * 2.0f;
                s->hk[set * H + i] = c;
                if (++k == N) k = 0;
            }
            j += m;
            if (j >= N) j -= N;
        }
    }
}

static void stage_free(stage_t *s) { free(s->h); free(s->d); free(s->hk); }

/* msd.hpp:228-247 */
static inline int add_sample(stage_t *s, cf sample) {
    int ret = 0;
    s->d[s->pos] = sample;
    s->d[s->pos + s->H] = sample;
    if (++s->pos == s->H) s->pos = 0;
    if (--s->isn == 0) { s->isn = s->m; ret = 1; }
    return ret;
}

/* msd.hpp:308-332, the portable folded FIR.  The reference unrolls by four but each of
 * real_sum / imag_sum is still one strictly sequential chain i = 0 .. half-1, then the
 * centre tap, so the plain loop below performs the identical float operations in order. */
static inline cf calculate_output(const stage_t *s) {
    size_t half = (s->H - 1) >> 1;
    size_t j = s->H - 1;
    float real_sum = 0.0f, imag_sum = 0.0f;
    const cf *d = s->d + s->pos;
    const float *h = s->h;
    size_t i = 0;
    for (; i < half; ++i, --j) {
        real_sum += (d[i].re + d[j].re) * h[i];
        imag_sum += (d[i].im + d[j].im) * h[i];
    }
    real_sum += d[i].re * h[i];
    imag_sum += d[i].im * h[i];
    cf r = { real_sum, imag_sum };
    return r;
}

/* msd.hpp:398-420, portable variant */
static inline cf calculate_output_translated(stage_t *s) {
    float real_sum = 0.0f, imag_sum = 0.0f;
    const cf *d = s->d + s->pos;
    const cf *h = s->hk + s->k * s->H;
    for (size_t i = 0; i < s->H; ++i) {
        real_sum += (d[i].re * h[i].re - d[i].im * h[i].im);
        imag_sum += (d[i].re * h[i].im + d[i].im * h[i].re);
    }
    if (++s->k == s->num_sets) s->k = 0;
    cf r = { real_sum, imag_sum };
    return r;
}

/* msd.hpp:54-68 */
sdrxo_msd *sdrxo_msd_create(const float *translator, int N, int n_stages, const int *m,
                            const int *ntaps, const float *taps, int use_ftfir) {
    sdrxo_msd *q = (sdrxo_msd *)calloc(1, sizeof(*q));
    q->n_stages = n_stages;
    q->st = (stage_t *)calloc((size_t)n_stages, sizeof(stage_t));
    q->m_total = 1;
    q->N = (translator && N > 0) ? (size_t)N : 0;
    q->translator = NULL;
    if (q->N) {
        q->translator = (cf *)malloc(sizeof(cf) * q->N);
        memcpy(q->translator, translator, sizeof(cf) * q->N);
    }
    q->trans_pos = 0;
    q->use_ftfir = use_ftfir;
    int pos = 0;
    for (int s = 0; s < n_stages; ++s) {
        if (s == 0 && q->N)
            stage_init(&q->st[s], (size_t)m[s], taps + pos, (size_t)ntaps[s], q->translator, q->N);
        else
            stage_init(&q->st[s], (size_t)m[s], taps + pos, (size_t)ntaps[s], NULL, 0);
        q->m_total *= (size_t)m[s];
        pos += ntaps[s];
    }
    return q;
}

sdrxo_msd *sdrxo_msd_clone(const sdrxo_msd *o) {
    sdrxo_msd *q = (sdrxo_msd *)malloc(sizeof(*q));
    *q = *o;
    q->st = (stage_t *)malloc(sizeof(stage_t) * (size_t)o->n_stages);
    for (int s = 0; s < o->n_stages; ++s) {
        const stage_t *a = &o->st[s];
        stage_t *b = &q->st[s];
        *b = *a;
        b->h = (float *)malloc(sizeof(float) * a->H); memcpy(b->h, a->h, sizeof(float) * a->H);
        b->d = (cf *)malloc(sizeof(cf) * 2 * a->H);   memcpy(b->d, a->d, sizeof(cf) * 2 * a->H);
        if (a->hk) {
            b->hk = (cf *)malloc(sizeof(cf) * a->num_sets * a->H);
            memcpy(b->hk, a->hk, sizeof(cf) * a->num_sets * a->H);
        }
    }
    if (o->translator) {
        q->translator = (cf *)malloc(sizeof(cf) * o->N);
        memcpy(q->translator, o->translator, sizeof(cf) * o->N);
    }
    return q;
}

void sdrxo_msd_destroy(sdrxo_msd *q) {
    if (!q) return;
    for (int s = 0; s < q->n_stages; ++s) stage_free(&q->st[s]);
    free(q->st); free(q->translator); free(q);
}

size_t sdrxo_msd_m(const sdrxo_msd *q) { return q->m_total; }

/* msd.hpp:75-159 */
size_t sdrxo_msd_decimate(sdrxo_msd *q, const float *in_f, size_t in_len, float *out_f) {
    const cf *in = (const cf *)in_f;
    cf *out = (cf *)out_f;
    size_t out_len = 0;
    cf sample;

    if (q->N == 0) {
        /* msd.hpp:79-101 no tuning required */
        for (size_t i = 0; i < in_len; ++i) {
            sample = in[i];
            int s = 0;
            while (s < q->n_stages) {
                if (add_sample(&q->st[s], sample)) sample = calculate_output(&q->st[s]);
                else break;
                ++s;
            }
            if (s == q->n_stages) { *(out++) = sample; out_len += 1; }
        }
    } else if (q->use_ftfir) {
        /* msd.hpp:104-131 */
        for (size_t i = 0; i < in_len; ++i) {
            sample = in[i];
            int s = 0;
            if (add_sample(&q->st[0], sample)) {
                sample = calculate_output_translated(&q->st[0]);
                ++s;
                while (s < q->n_stages) {
                    if (add_sample(&q->st[s], sample)) sample = calculate_output(&q->st[s]);
                    else break;
                    ++s;
                }
            }
            if (s == q->n_stages) { *(out++) = sample; out_len += 1; }
        }
    } else {
        /* msd.hpp:133-154 */
        for (size_t i = 0; i < in_len; ++i) {
            cf t = q->translator[q->trans_pos];
            /* std::complex<float> operator*: (ac - bd) + (ad + bc)i */
            sample.re = in[i].re * t.re - in[i].im * t.im;
            sample.im = in[i].re * t.im + in[i].im * t.re;
            if (++q->trans_pos == q->N) q->trans_pos = 0;
            int s = 0;
            while (s < q->n_stages) {
                if (add_sample(&q->st[s], sample)) sample = calculate_output(&q->st[s]);
                else break;
                ++s;
            }
            if (s == q->n_stages) { *(out++) = sample; out_len += 1; }
        }
    }
    return out_len;
}

/* ------------------------------------------------------------------ AGC / LfAGC (agc.hpp) */

/* agc.hpp:30-36 */
void sdrxo_agc_init(sdrxo_agc *a, float attack, float decay, float reference, float max_gain, float alpha) {
    a->attack = attack; a->decay = decay; a->reference = reference; a->max_gain = max_gain; a->alpha = alpha;
    a->gain = 1.0f; a->power = reference;
    a->min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
}
/* agc.hpp:41-44 */
void sdrxo_agc_set_max_gain(sdrxo_agc *a, float max_gain) {
    a->max_gain = max_gain;
    a->min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
}
/* agc.hpp:48-69 */
static inline cf agc_adjust1(sdrxo_agc *a, cf sample) {
    cf y = { sample.re * a->gain, sample.im * a->gain };
    a->power += a->alpha * ((y.re * y.re + y.im * y.im) - a->power);
    float error = a->reference - a->power;
    if (error > 0.0f) a->gain += a->decay * error;
    else              a->gain += a->attack * error;
    if (a->gain < a->min_gain) a->gain = a->min_gain;
    else if (a->gain > a->max_gain) a->gain = a->max_gain;
    return y;
}
void sdrxo_agc_adjust(sdrxo_agc *a, const float *in, unsigned n, float *out) {
    for (unsigned i = 0; i < n; ++i) {
        cf s = { in[2 * i], in[2 * i + 1] };
        cf y = agc_adjust1(a, s);
        out[2 * i] = y.re; out[2 * i + 1] = y.im;
    }
}
/* agc.hpp:86-91 */
void sdrxo_lfagc_init(sdrxo_lfagc *a, float attack, float decay, float reference, float max_gain, float alpha) {
    a->attack = attack; a->decay = decay; a->reference = reference; a->max_gain = max_gain; a->alpha = alpha;
    a->gain = 1.0f; a->power = reference * reference; a->active = 0;
    a->min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
}
void sdrxo_lfagc_set_max_gain(sdrxo_lfagc *a, float max_gain) {
    a->max_gain = max_gain;
    a->min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
}
/* agc.hpp:106-131 */
static inline float lfagc_adjust1(sdrxo_lfagc *a, float sample) {
    float y = sample * a->gain;
    a->power += a->alpha * (y * y - a->power);
    float error = a->reference - sqrtf(a->power);
    if (error > 0.0f) a->gain += a->decay * error;
    else              a->gain += a->attack * error;
    if (a->gain < a->min_gain) a->gain = a->min_gain;
    else if (a->gain > a->max_gain) a->gain = a->max_gain;
    return a->active ? y : sample;
}
void sdrxo_lfagc_adjust(sdrxo_lfagc *a, const float *in, unsigned n, float *out) {
    for (unsigned i = 0; i < n; ++i) out[i] = lfagc_adjust1(a, in[i]);
}

/* ------------------------------------------------------------------ FIR family (fir.hpp) */

struct sdrxo_fir {
    int      kind;
    unsigned size, pos;
    float   *c, *c_adj;
    float   *buf_a, *buf_b;   /* FIR: buf_a; FIR2: bufr=buf_a, bufl=buf_b; FIR3c: re=buf_a, im=buf_b */
    float    gain;
};

sdrxo_fir *sdrxo_fir_create(int kind, const float *c, int n) {
    sdrxo_fir *f = (sdrxo_fir *)calloc(1, sizeof(*f));
    f->kind = kind; f->size = (unsigned)n; f->pos = 0; f->gain = 0.0f;
    f->c = (float *)malloc(sizeof(float) * (size_t)n);
    f->c_adj = (float *)malloc(sizeof(float) * (size_t)n);
    memcpy(f->c, c, sizeof(float) * (size_t)n);
    memcpy(f->c_adj, c, sizeof(float) * (size_t)n);
    f->buf_a = (float *)calloc((size_t)n, sizeof(float));
    f->buf_b = (float *)calloc((size_t)n, sizeof(float));
    return f;
}
void sdrxo_fir_destroy(sdrxo_fir *f) {
    if (!f) return;
    free(f->c); free(f->c_adj); free(f->buf_a); free(f->buf_b); free(f);
}
/* fir.hpp:50-60, :101-111, :160-170 */
void sdrxo_fir_set_gain(sdrxo_fir *f, float gain_db) {
    f->gain = gain_db;
    for (unsigned i = 0; i < f->size; ++i) f->c_adj[i] = f->c[i] * powf(10.0f, f->gain / 20.0f);
}
void sdrxo_fir_filter(sdrxo_fir *f, const float *in, unsigned in_len, float *out) {
    if (f->kind == 1) {
        /* fir.hpp:85-99 */
        for (unsigned n = 0; n < in_len; ++n) {
            f->buf_a[f->pos] = in[n];
            if (++f->pos == f->size) f->pos = 0;
            float acc = 0.0f;
            for (unsigned i = 0; i < f->size; ++i) {
                acc += f->c_adj[i] * f->buf_a[f->pos];
                if (++f->pos == f->size) f->pos = 0;
            }
            out[n] = acc;
        }
    } else if (f->kind == 2) {
        /* fir.hpp:137-157 */
        for (unsigned n = 0; n < in_len; n += 2) {
            f->buf_a[f->pos] = in[n];
            f->buf_b[f->pos] = in[n + 1];
            if (++f->pos == f->size) f->pos = 0;
            float r = 0.0f, l = 0.0f;
            for (unsigned i = 0; i < f->size; ++i) {
                r += f->c_adj[i] * f->buf_a[f->pos];
                l += f->c_adj[i] * f->buf_b[f->pos];
                if (++f->pos == f->size) f->pos = 0;
            }
            out[n] = r; out[n + 1] = l;
        }
    } else {
        /* fir.hpp:34-48 with T = std::complex<float>: *out += float * complex */
        for (unsigned n = 0; n < in_len; ++n) {
            f->buf_a[f->pos] = in[2 * n];
            f->buf_b[f->pos] = in[2 * n + 1];
            if (++f->pos == f->size) f->pos = 0;
            float re = 0.0f, im = 0.0f;
            for (unsigned i = 0; i < f->size; ++i) {
                re += f->buf_a[f->pos] * f->c_adj[i];
                im += f->buf_b[f->pos] * f->c_adj[i];
                if (++f->pos == f->size) f->pos = 0;
            }
            out[2 * n] = re; out[2 * n + 1] = im;
        }
    }
}

/* ------------------------------------------------------------------ receiver chain */

enum { SQL_CLOSED = 0, SQL_OPEN = 1 }; /* sdrx.cpp:149 */

typedef struct {
    sdrxo_msd  *msd;
    sdrxo_agc   agc;
    sdrxo_lfagc agc_lf;
    float sql_level;
    int   sql_state, sql_state_prev, pos;
    int   mod;                      /* 0 AM, 1 FM: Channel::mod (sdrx.cpp:230) */
    cf    prev;                     /* Demod::prev_sample (sdrx.cpp:202), starts at (0, 0) */
} chan_t;

struct sdrxo_rx {
    int        C;
    chan_t    *ch;
    sdrxo_fir *audio_filter;
    float      window[FFT_SIZE];
    float      passband_shape[FFT_SIZE], ramp_up[CH_IQ_BUF_SIZE], ramp_down[CH_IQ_BUF_SIZE];
    cf        *iq_in; unsigned iq_len, iq_cap;
    cf        *chunk;
};

/* sdrx.cpp:1680-1714, :880-881, :898-900 */
sdrxo_rx *sdrxo_rx_create(const sdrxo_rx_config *cfg) {
    sdrxo_rx *rx = (sdrxo_rx *)calloc(1, sizeof(*rx));
    rx->C = cfg->n_channels;
    rx->ch = (chan_t *)calloc((size_t)rx->C, sizeof(chan_t));
    for (int c = 0; c < rx->C; ++c) {
        chan_t *ch = &rx->ch[c];
        const float *tr = (cfg->has_translator && cfg->has_translator[c])
                              ? cfg->translators + (size_t)c * (size_t)cfg->N * 2 : NULL;
        ch->msd = sdrxo_msd_create(tr, tr ? cfg->N : 0, cfg->n_stages, cfg->m, cfg->ntaps, cfg->taps,
                                   cfg->use_ftfir);
        /* Channel() default-constructs AGC/LfAGC (agc.hpp:30, :86), then the setters of
         * sdrx.cpp:1702-1710 run -- power_ keeps the constructor's reference. */
        sdrxo_agc_init(&ch->agc, 10.0f, 0.01f, 0.25f, 200.0f, 0.01f);
        ch->agc.reference = 1.0f; ch->agc.attack = 1.0f; ch->agc.decay = 0.01f;
        sdrxo_agc_set_max_gain(&ch->agc, 300);
        sdrxo_lfagc_init(&ch->agc_lf, 10.0f, 0.01f, 0.25f, 200.0f, 0.005f);
        ch->agc_lf.reference = 1.0f; ch->agc_lf.attack = 1.0f; ch->agc_lf.decay = 0.01f;
        if (cfg->use_lf_agc) ch->agc_lf.active = 1;
        ch->sql_level = cfg->sql_level ? cfg->sql_level[c] : 9.0f;
        ch->sql_state = SQL_CLOSED; ch->sql_state_prev = SQL_CLOSED;
        ch->pos = cfg->pos ? cfg->pos[c] : sdrxo_audio_pos((unsigned)c, (unsigned)rx->C);
    }
    rx->audio_filter = sdrxo_fir_create(2, cfg->fir2_taps, cfg->fir2_ntaps);
    sdrxo_fir_set_gain(rx->audio_filter, cfg->lf_gain_db);
    sdrxo_window(rx->window);
    memcpy(rx->passband_shape, cfg->passband_shape, sizeof(rx->passband_shape));
    memcpy(rx->ramp_up, cfg->ramp_up, sizeof(rx->ramp_up));
    memcpy(rx->ramp_down, cfg->ramp_down, sizeof(rx->ramp_down));
    rx->chunk = (cf *)calloc((size_t)rx->C * CH_IQ_BUF_SIZE, sizeof(cf));
    return rx;
}

void sdrxo_rx_destroy(sdrxo_rx *rx) {
    if (!rx) return;
    for (int c = 0; c < rx->C; ++c) sdrxo_msd_destroy(rx->ch[c].msd);
    free(rx->ch); sdrxo_fir_destroy(rx->audio_filter); free(rx->iq_in); free(rx->chunk); free(rx);
}

float sdrxo_rx_ingest(sdrxo_rx *rx, const void *raw, int fmt, unsigned n) {
    if (n > rx->iq_cap) { rx->iq_in = (cf *)realloc(rx->iq_in, sizeof(cf) * n); rx->iq_cap = n; }
    rx->iq_len = n;
    cf *iq = rx->iq_in;
    if (fmt == 0) {
        /* rtl_dev.cpp:316-327 */
        const unsigned char *data = (const unsigned char *)raw;
        for (unsigned i = 0; i < n; ++i) {
            iq[i].re = ((float)data[2 * i]) / 127.5f - 1.0f;
            iq[i].im = ((float)data[2 * i + 1]) / 127.5f - 1.0f;
        }
    } else if (fmt == 1) {
        /* libairspy INT16_IQ is float*32768 (airspy.c:58-62, :304); the inverse is exact */
        const int16_t *data = (const int16_t *)raw;
        for (unsigned i = 0; i < n; ++i) {
            iq[i].re = (float)data[2 * i] * (1.0f / 32768.0f);
            iq[i].im = (float)data[2 * i + 1] * (1.0f / 32768.0f);
        }
    } else {
        /* airspy_dev.cpp:355-357 */
        memcpy(iq, raw, sizeof(cf) * n);
    }
    /* rtl_dev.cpp:329-341 == airspy_dev.cpp:394-405 (portable variant) */
    float pwr_rms = 0.0f;
    for (unsigned i = 0; i < n; i++) {
        float ampl_squared = iq[i].re * iq[i].re + iq[i].im * iq[i].im;
        pwr_rms += ampl_squared;
    }
    pwr_rms = pwr_rms / n;
    return 10 * log10f(pwr_rms) - 3.0f;
}

/* sdrx.cpp:342-345 */
void sdrxo_rx_channelize(sdrxo_rx *rx, float *ch_iq_out) {
    for (int c = 0; c < rx->C; ++c)
        sdrxo_msd_decimate(rx->ch[c].msd, (const float *)rx->iq_in, rx->iq_len,
                           (float *)(rx->chunk + (size_t)c * CH_IQ_BUF_SIZE));
    if (ch_iq_out) memcpy(ch_iq_out, rx->chunk, sizeof(cf) * (size_t)rx->C * CH_IQ_BUF_SIZE);
}

/* sdrx.cpp:471-497 */
static inline void pan_mix(float *mix, unsigned i, int pos, float s) {
    switch (pos) {
        case -2: mix[i * 2] += 0.8f * s; mix[i * 2 + 1] += 0.2f * s; break;
        case -1: mix[i * 2] += 0.6f * s; mix[i * 2 + 1] += 0.4f * s; break;
        case 1:  mix[i * 2] += 0.4f * s; mix[i * 2 + 1] += 0.6f * s; break;
        case 2:  mix[i * 2] += 0.2f * s; mix[i * 2 + 1] += 0.8f * s; break;
        default: mix[i * 2] += 0.5f * s; mix[i * 2 + 1] += 0.5f * s; break;
    }
}

static inline float norm2(cf z) { return z.re * z.re + z.im * z.im; }

/* Demod::demod (sdrx.cpp:178-200).  AM: the envelope.  FM (flagged EXPERIMENTAL, sdrx.cpp:1197): the sample is
 * normalised (std::complex<float> / float divides both parts), then the angle from the previous normalised sample
 * to this one, atan2(q*pr - i*pi, i*pr + q*pi); prev_sample only moves when demod() is called, i.e. while the
 * squelch is open (sdrx.cpp:459-461) -- the ramp-down block uses the envelope whatever the modulation (:500). */
static inline float demod1(chan_t *ch, cf a) {
    const float mag = hypotf(a.re, a.im);
    if (ch->mod != 1) return mag;
    const float i = a.re / mag, q = a.im / mag;
    const float s = atan2f(q * ch->prev.re - i * ch->prev.im, i * ch->prev.re + q * ch->prev.im);
    ch->prev.re = i; ch->prev.im = q;
    return s;
}

void sdrxo_rx_set_modulation(sdrxo_rx *rx, const int *mod) {
    for (int c = 0; c < rx->C; ++c) rx->ch[c].mod = mod ? mod[c] : 0;
}

/* sdrx.cpp:452-573 then :646-656 */
void sdrxo_rx_demod(sdrxo_rx *rx, float *audio, uint8_t *sql_state, float *snr_out,
                    float *mix_pre, float *mix, int16_t *pcm) {
    float   audio_buffer_float[CH_IQ_BUF_SIZE * 2];
    cf      fft_in[FFT_SIZE], fft_out[FFT_SIZE];

    memset(audio_buffer_float, 0, sizeof(audio_buffer_float));
    if (audio) memset(audio, 0, sizeof(float) * (size_t)rx->C * CH_IQ_BUF_SIZE);
    unsigned j = 0;
    for (int c = 0; c < rx->C; ++c) {
        chan_t *ch = &rx->ch[c];
        for (unsigned i = 0; i < CH_IQ_BUF_SIZE; ++i) {
            cf a = agc_adjust1(&ch->agc, rx->chunk[j]);                 /* :457 */
            if (ch->sql_state == SQL_OPEN) {
                float s = demod1(ch, a);                                /* :461, Demod :178-200 */
                s = lfagc_adjust1(&ch->agc_lf, s);                      /* :463 */
                if (ch->sql_state_prev == SQL_CLOSED) s = rx->ramp_up[i] * s;   /* :466-468 */
                pan_mix(audio_buffer_float, i, ch->pos, s);
                if (audio) audio[(size_t)c * CH_IQ_BUF_SIZE + i] = s;
            } else if (ch->sql_state_prev == SQL_OPEN) {
                float s = hypotf(a.re, a.im);                           /* :500 */
                s = lfagc_adjust1(&ch->agc_lf, s);
                s = rx->ramp_down[i] * s;
                pan_mix(audio_buffer_float, i, ch->pos, s);
                if (audio) audio[(size_t)c * CH_IQ_BUF_SIZE + i] = s;
            }
            fft_in[i].re = rx->chunk[j].re * rx->window[i];             /* :533 */
            fft_in[i].im = rx->chunk[j].im * rx->window[i];
            ++j;
        }
        ch->sql_state_prev = ch->sql_state;                             /* :536 */

        sdrx_oracle_fft512((const float *)fft_in, (float *)fft_out);    /* :538 */

        float sig_level = 0.0f;                                         /* :542-550 */
        for (unsigned i = 3; i < 91; i++) {
            sig_level += norm2(fft_out[i]);
            sig_level += norm2(fft_out[FFT_SIZE - i]);
        }
        sig_level /= 176;

        float ref_level_hi = 0.0f, ref_level_lo = 0.0f;                 /* :552-562 */
        for (unsigned i = 112; i < 157; i++) {
            cf hi = { fft_out[i].re * rx->passband_shape[i], fft_out[i].im * rx->passband_shape[i] };
            cf lo = { fft_out[FFT_SIZE - i].re * rx->passband_shape[FFT_SIZE - i],
                      fft_out[FFT_SIZE - i].im * rx->passband_shape[FFT_SIZE - i] };
            ref_level_hi += norm2(hi);
            ref_level_lo += norm2(lo);
        }
        ref_level_hi /= 45;
        ref_level_lo /= 45;
        float noise_level = (ref_level_hi + ref_level_lo) / 2;

        float snr = 10 * log10f(sig_level / noise_level);               /* :566 */

        if (snr > ch->sql_level + 3 || ch->sql_level == 0.0f) ch->sql_state = SQL_OPEN;   /* :569-573 */
        else if (snr < ch->sql_level) ch->sql_state = SQL_CLOSED;

        if (sql_state) sql_state[c] = (uint8_t)ch->sql_state;
        if (snr_out) snr_out[c] = snr;
    }

    if (mix_pre) memcpy(mix_pre, audio_buffer_float, sizeof(audio_buffer_float));
    sdrxo_fir_filter(rx->audio_filter, audio_buffer_float, CH_IQ_BUF_SIZE * 2, audio_buffer_float); /* :646 */
    if (mix) memcpy(mix, audio_buffer_float, sizeof(audio_buffer_float));
    if (pcm) {
        for (unsigned i = 0; i < CH_IQ_BUF_SIZE * 2; ++i) {             /* :649-656 */
            int16_t s;
            if (audio_buffer_float[i] > 1.0f) s = 32767;
            else if (audio_buffer_float[i] < -1.0f) s = -32767;
            else s = (int16_t)(audio_buffer_float[i] * 32767.0f);
            pcm[i] = s;
        }
    }
}

void sdrxo_rx_gains(sdrxo_rx *rx, float *agc_gain, float *lf_gain) {
    for (int c = 0; c < rx->C; ++c) {
        if (agc_gain) agc_gain[c] = rx->ch[c].agc.gain;
        if (lf_gain) lf_gain[c] = rx->ch[c].agc_lf.gain;
    }
}

/* TEST INFRASTRUCTURE -- CPU restatement of the Airspy real -> IQ front end, the step in
 * front of the sdrx hot path that libairspy runs on the host for AIRSPY_SAMPLE_FLOAT32_IQ.  Only tests/,
 * __graft_entry__.smoke() and bench.py's CPU legs may load this; the product never does.
 *
 * Follows (paths under /root/reference/libairspy/libairspy/src):
 *   airspy.c:323-338   unpack_samples        12-bit samples, 8 per three 32-bit words
 *   airspy.c:311-321   convert_samples_float (raw - 2048) * SAMPLE_SCALE, SAMPLE_SCALE = 1/2048 (airspy.c:58-62)
 *   airspy.c:393-396   the FLOAT32_IQ case: convert, iqconverter_float_process, sample_count /= 2
 *   iqconverter_float.c:74-96  create: len = 47/2+1 = 24, hbc = centre tap, fir_kernel[i] = hb_kernel[2i]
 *   iqconverter_float.c:455-468  remove_dc: s -= avg; avg += 0.01f * s      (one running mean over ALL samples)
 *   iqconverter_float.c:470-502  translate_fs_4: s[4i] = -s, s[4i+1] = -s*hbc, s[4i+3] = s*hbc
 *   iqconverter_float.c:191-230  fir_interleaved_24 on the even samples: sum of k[i]*(q[i]+q[23-i]), i = 0..11,
 *                                added left to right, q[0] the newest even sample
 *   iqconverter_float.c:430-451  delay_interleaved on the odd samples: a 12-deep delay line
 * Pinned bit-exact against oracle/_ref/libairspy_ref.so (the reference's own iqconverter_float.c compiled with
 * the oracle's portable flags) by tests/test_airspy_fe.py.  The ring buffers of the reference are replaced by
 * plain sliding histories; the arithmetic and its order are the reference's. */
#include "airspy_fe_oracle.h"

#include <stdlib.h>
#include <string.h>

struct sdrxo_airspy_fe {
    int    len;          /* 24 */
    float  hbc;
    float *k;            /* [len] every second tap */
    float  avg;
    float *even;         /* last len-1 rotated even samples, oldest first */
    float *odd;          /* last len/2 rotated+scaled odd samples, oldest first */
};

sdrxo_airspy_fe *sdrxo_airspy_fe_create(const float *hb_kernel, int hb_len) {
    sdrxo_airspy_fe *f = (sdrxo_airspy_fe *)calloc(1, sizeof(*f));
    f->len = hb_len / 2 + 1;
    f->hbc = hb_kernel[hb_len / 2];
    f->k = (float *)malloc(sizeof(float) * (size_t)f->len);
    for (int i = 0, j = 0; i < f->len; i++, j += 2) f->k[i] = hb_kernel[j];
    f->even = (float *)calloc((size_t)f->len, sizeof(float));
    f->odd = (float *)calloc((size_t)f->len, sizeof(float));
    f->avg = 0.0f;
    return f;
}

void sdrxo_airspy_fe_destroy(sdrxo_airspy_fe *f) {
    if (!f) return;
    free(f->k); free(f->even); free(f->odd); free(f);
}

void sdrxo_airspy_fe_reset(sdrxo_airspy_fe *f) {
    f->avg = 0.0f;
    memset(f->even, 0, sizeof(float) * (size_t)f->len);
    memset(f->odd, 0, sizeof(float) * (size_t)f->len);
}

float sdrxo_airspy_fe_avg(const sdrxo_airspy_fe *f) { return f->avg; }

void sdrxo_airspy_unpack12(const uint32_t *in, uint16_t *out, int n) {
    /* 96 bits hold eight samples, most significant first across the three words */
    for (int i = 0, j = 0; j < n; i += 3, j += 8) {
        const uint32_t a = in[i], b = in[i + 1], c = in[i + 2];
        out[j + 0] = (uint16_t)((a >> 20) & 0xfff);
        out[j + 1] = (uint16_t)((a >> 8) & 0xfff);
        out[j + 2] = (uint16_t)(((a & 0xff) << 4) | ((b >> 28) & 0xf));
        out[j + 3] = (uint16_t)((b >> 16) & 0xfff);
        out[j + 4] = (uint16_t)((b >> 4) & 0xfff);
        out[j + 5] = (uint16_t)(((b & 0xf) << 8) | ((c >> 24) & 0xff));
        out[j + 6] = (uint16_t)((c >> 12) & 0xfff);
        out[j + 7] = (uint16_t)(c & 0xfff);
    }
}

void sdrxo_airspy_fe_process(sdrxo_airspy_fe *f, const uint16_t *raw, int n, float *iq_out) {
    const int L = f->len, half = L >> 1;
    const int npair = n / 2;
    /* working rows: [history | this call] */
    float *ev = (float *)malloc(sizeof(float) * (size_t)(L - 1 + npair));
    float *od = (float *)malloc(sizeof(float) * (size_t)(half + npair));
    memcpy(ev, f->even, sizeof(float) * (size_t)(L - 1));
    memcpy(od, f->odd, sizeof(float) * (size_t)half);
    float avg = f->avg;
    for (int p = 0; p < npair; ++p) {
        /* convert_samples_float + remove_dc, sample by sample in stream order */
        float s0 = (float)((int)raw[2 * p] - 2048) * (1.0f / 2048.0f);
        s0 -= avg;
        avg += 0.01f * s0;
        float s1 = (float)((int)raw[2 * p + 1] - 2048) * (1.0f / 2048.0f);
        s1 -= avg;
        avg += 0.01f * s1;
        /* translate_fs_4: pair p is (4i, 4i+1) when p is even, (4i+2, 4i+3) when odd */
        if ((p & 1) == 0) { s0 = -s0; s1 = -s1 * f->hbc; }
        else              { s1 = s1 * f->hbc; }
        ev[L - 1 + p] = s0;
        od[half + p] = s1;
    }
    for (int p = 0; p < npair; ++p) {
        const float *q = ev + (L - 1) + p;          /* q[-i] is the i-th older even sample */
        float acc = f->k[0] * (q[0] + q[-(L - 1)]);
        for (int i = 1; i < half; ++i) acc = acc + f->k[i] * (q[-i] + q[-(L - 1 - i)]);
        iq_out[2 * p] = acc;
        iq_out[2 * p + 1] = od[p];                   /* the sample that entered the delay line `half` pairs ago */
    }
    memcpy(f->even, ev + npair, sizeof(float) * (size_t)(L - 1));
    memcpy(f->odd, od + npair, sizeof(float) * (size_t)half);
    f->avg = avg;
    free(ev); free(od);
}

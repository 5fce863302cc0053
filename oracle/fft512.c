/*
 * TEST INFRASTRUCTURE (oracle) -- see fft512.h.
 * Build with -O2 -ffp-contract=off so every multiply and add rounds on its own.
 */
#include "fft512.h"
#include <math.h>

#define NFFT 512
#define LOG2N 9

static float tw_re[NFFT / 2];
static float tw_im[NFFT / 2];
static unsigned short rev[NFFT];
static int ready = 0;

static void init_tables(void) {
    /* Twiddles exp(-2*pi*i*k/512): evaluated in double, rounded once to float. */
    for (int k = 0; k < NFFT / 2; ++k) {
        double a = -2.0 * M_PI * (double)k / (double)NFFT;
        tw_re[k] = (float)cos(a);
        tw_im[k] = (float)sin(a);
    }
    for (int i = 0; i < NFFT; ++i) {
        unsigned r = 0;
        for (int b = 0; b < LOG2N; ++b)
            if (i & (1 << b)) r |= 1u << (LOG2N - 1 - b);
        rev[i] = (unsigned short)r;
    }
    ready = 1;
}

void sdrx_oracle_fft512(const float *in, float *out) {
    if (!ready) init_tables();

    for (int i = 0; i < NFFT; ++i) {
        out[2 * rev[i] + 0] = in[2 * i + 0];
        out[2 * rev[i] + 1] = in[2 * i + 1];
    }

    for (int len = 2; len <= NFFT; len <<= 1) {
        int half = len >> 1;
        int step = NFFT / len;
        for (int base = 0; base < NFFT; base += len) {
            for (int j = 0; j < half; ++j) {
                float wr = tw_re[j * step];
                float wi = tw_im[j * step];
                float *a = out + 2 * (base + j);
                float *b = out + 2 * (base + j + half);
                float tr = wr * b[0] - wi * b[1];
                float ti = wr * b[1] + wi * b[0];
                b[0] = a[0] - tr;
                b[1] = a[1] - ti;
                a[0] = a[0] + tr;
                a[1] = a[1] + ti;
            }
        }
    }
}

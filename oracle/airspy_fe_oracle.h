/* TEST INFRASTRUCTURE (see airspy_fe_oracle.c). */
#ifndef SDRX_AIRSPY_FE_ORACLE_H
#define SDRX_AIRSPY_FE_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef struct sdrxo_airspy_fe sdrxo_airspy_fe;
sdrxo_airspy_fe *sdrxo_airspy_fe_create(const float *hb_kernel, int hb_len);
void sdrxo_airspy_fe_destroy(sdrxo_airspy_fe *f);
void sdrxo_airspy_fe_reset(sdrxo_airspy_fe *f);
/* n 12-bit samples packed 8 per three 32-bit words -> one per uint16 (n % 8 == 0) */
void sdrxo_airspy_unpack12(const uint32_t *in, uint16_t *out, int n);
/* n real ADC samples (one per uint16, 0..4095; n % 4 == 0) -> n/2 interleaved IQ pairs */
void sdrxo_airspy_fe_process(sdrxo_airspy_fe *f, const uint16_t *raw, int n, float *iq_out);
float sdrxo_airspy_fe_avg(const sdrxo_airspy_fe *f);
#ifdef __cplusplus
}
#endif
#endif

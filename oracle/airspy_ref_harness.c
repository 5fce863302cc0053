/* TEST INFRASTRUCTURE -- thin C harness around the reference's own libairspy IQ converter, compiled from where it
 * lies (/root/reference/libairspy/libairspy/src/iqconverter_float.c + filters.h) by oracle/Makefile into
 * oracle/_ref/libairspy_ref.so.  Nothing of the reference is copied: this file only forwards.
 * Used to pin oracle/airspy_fe_oracle.c (the Airspy real->IQ front end). */
#include <string.h>

#include "filters.h"              /* HB_KERNEL_FLOAT[47], libairspy/src/filters.h:28-30 */
#include "iqconverter_float.h"

int sdrxref_hb_kernel_len(void) { return HB_KERNEL_FLOAT_LEN; }
void sdrxref_hb_kernel(float *out) { memcpy(out, HB_KERNEL_FLOAT, sizeof(float) * HB_KERNEL_FLOAT_LEN); }

/* airspy.c:~500  device->cnv_f = iqconverter_float_create(HB_KERNEL_FLOAT, HB_KERNEL_FLOAT_LEN) */
void *sdrxref_iqconv_create(void) { return iqconverter_float_create(HB_KERNEL_FLOAT, HB_KERNEL_FLOAT_LEN); }
void sdrxref_iqconv_free(void *c) { iqconverter_float_free((iqconverter_float_t *)c); }
void sdrxref_iqconv_reset(void *c) { iqconverter_float_reset((iqconverter_float_t *)c); }
/* in place: len real samples in, len/2 interleaved IQ pairs out (airspy.c:393-396) */
void sdrxref_iqconv_process(void *c, float *samples, int len) { iqconverter_float_process((iqconverter_float_t *)c, samples, len); }

/* Declarations-only stand-in for <fftw3.h> (single-precision interface sdrx.cpp uses, sdrx.cpp:887-890, :538, :942-945). */
#ifndef STUB_FFTW3_H
#define STUB_FFTW3_H
#ifdef __cplusplus
extern "C" {
#endif
#define FFTW_FORWARD (-1)
#define FFTW_BACKWARD (+1)
#define FFTW_MEASURE (0U)
#define FFTW_ESTIMATE (1U << 6)
typedef float fftwf_complex[2];
typedef struct fftwf_plan_s *fftwf_plan;
fftwf_plan fftwf_plan_dft_1d(int n, fftwf_complex *in, fftwf_complex *out, int sign, unsigned flags);
void fftwf_execute(const fftwf_plan p);
void fftwf_destroy_plan(fftwf_plan p);
void fftwf_cleanup(void);
void *fftwf_malloc(size_t n);
void fftwf_free(void *p);
#ifdef __cplusplus
}
#endif
#endif

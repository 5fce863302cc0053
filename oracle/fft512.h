/*
 * TEST INFRASTRUCTURE (oracle) -- not part of the product path.
 *
 * Plain FP32 radix-2 decimation-in-time forward DFT, 512 points, unnormalised,
 * out of place.  It stands in for the reference's call to FFTW3f
 *   plan:    /root/reference/src/sdrx.cpp:887-890  fftwf_plan_dft_1d(512, FORWARD, ESTIMATE)
 *   execute: /root/reference/src/sdrx.cpp:538
 * FFTW3f is an un-vendored system dependency with no pinned version
 * (/root/reference/CMakeLists.txt:50), so the squelch FFT boundary is
 * "parity unpinned": this file DEFINES the oracle's FFT, and is cross-checked
 * against numpy's double precision FFT in tests/test_oracle.py.
 */
#ifndef SDRX_ORACLE_FFT512_H
#define SDRX_ORACLE_FFT512_H

#ifdef __cplusplus
extern "C" {
#endif

/* in/out: 512 interleaved (re,im) float pairs. in != out. */
void sdrx_oracle_fft512(const float *in, float *out);

#ifdef __cplusplus
}
#endif
#endif

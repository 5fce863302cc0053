// sdrx_b200 drop-in for sdrx's src/iqsample.hpp (reference iqsample.hpp:26): the sample type is unchanged,
// interleaved (re, im) FP32 -- exactly the float pairs the C ABI of ../sdrx_b200.h takes.
#ifndef SDRX_B200_IQSAMPLE_HPP
#define SDRX_B200_IQSAMPLE_HPP

#include <complex>

typedef std::complex<float> iqsample_t;

#endif

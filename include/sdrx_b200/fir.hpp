// sdrx_b200 drop-in for sdrx's src/fir.hpp: FIR3<T> (fir.hpp:27-74), FIR (fir.hpp:78-127) and FIR2
// (fir.hpp:130-182) with the reference's surface -- ctor from a coefficient vector, filter(in, in_len, out)
// with in == out allowed, setGain(dB), gain() -- backed by the CUDA library (sdrx_fir_* in ../sdrx_b200.h).
// in_len follows the reference: floats for FIR and FIR2 (so 2 x frames for the stereo FIR2), samples for
// FIR3<T>.  The ring-buffer delay line of the reference lives on the device between calls.
#ifndef SDRX_B200_FIR_HPP
#define SDRX_B200_FIR_HPP

#include <complex>
#include <cstdio>
#include <type_traits>
#include <vector>

#include <sdrx_b200.h>       // -I<repo>/include (the headers themselves may be copied next to sdrx.cpp)

namespace sdrx_b200_detail {
class FirHandle {          // shared by the three classes; a copy carries taps, gain and the delay line, like the reference's implicit copy (fir.hpp:69-74)
public:
    FirHandle() = default;
    FirHandle(int kind, const std::vector<float> &c, int device = 0) : kind_(kind), c_(c), device_(device) { create_(); }
    FirHandle(const FirHandle &o) : kind_(o.kind_), c_(o.c_), device_(o.device_), gain_(o.gain_) {
        if (!o.h_) return;
        int rc = sdrx_fir_clone(o.h_, &h_);
        if (rc) std::fprintf(stderr, "sdrx_b200: sdrx_fir_clone failed (%d): %s\n", rc, sdrx_last_error());
    }
    FirHandle &operator=(FirHandle o) { swap_(o); return *this; }
    ~FirHandle() { sdrx_fir_destroy(h_); }
    void filter(const float *in, unsigned in_len, float *out) {
        if (!h_ || in_len == 0) return;
        int rc = sdrx_fir_filter(h_, in, in_len, out);
        if (rc) std::fprintf(stderr, "sdrx_b200: sdrx_fir_filter failed (%d): %s\n", rc, sdrx_last_error());
    }
    void setGain(float gain) { gain_ = gain; if (h_) sdrx_fir_set_gain(h_, gain); }
    float gain() const { return gain_; }
private:
    int kind_ = 1;
    std::vector<float> c_;
    int device_ = 0;
    float gain_ = 0.0f;
    sdrx_fir *h_ = nullptr;
    void create_() {
        if (c_.empty()) return;
        int rc = sdrx_fir_create(kind_, c_.data(), (int)c_.size(), device_, &h_);
        if (rc) std::fprintf(stderr, "sdrx_b200: sdrx_fir_create failed (%d): %s\n", rc, sdrx_last_error());
    }
    void swap_(FirHandle &o) {
        std::swap(kind_, o.kind_); c_.swap(o.c_); std::swap(device_, o.device_); std::swap(gain_, o.gain_); std::swap(h_, o.h_);
    }
};
}  // namespace sdrx_b200_detail

// Real (T = float) or complex (T = std::complex<float>) samples.
template <typename T = float>
class FIR3 {
    static_assert(std::is_same<T, float>::value || std::is_same<T, std::complex<float>>::value,
                  "FIR3<T>: T is float or std::complex<float>");
public:
    FIR3(const std::vector<float> c = std::vector<float>())
        : f_(std::is_same<T, float>::value ? 1 : 3, c) {}
    void filter(const T *in, unsigned in_len, T *out) {
        f_.filter(reinterpret_cast<const float *>(in), in_len, reinterpret_cast<float *>(out));
    }
    void setGain(float gain) { f_.setGain(gain); }
    float gain() const { return f_.gain(); }
private:
    sdrx_b200_detail::FirHandle f_;
};

class FIR {
public:
    FIR(const std::vector<float> c) : f_(1, c) {}
    void filter(const float *in, unsigned in_len, float *out) { f_.filter(in, in_len, out); }
    void setGain(float gain) { f_.setGain(gain); }
    float gain() const { return f_.gain(); }
private:
    sdrx_b200_detail::FirHandle f_;
};

// Stereo version: ALSA interleaved frames, both lanes through the same taps.
class FIR2 {
public:
    FIR2() = default;
    FIR2(const std::vector<float> c) : f_(2, c) {}
    void filter(const float *in, unsigned in_len, float *out) { f_.filter(in, in_len, out); }
    void setGain(float gain) { f_.setGain(gain); }
    float gain() const { return f_.gain(); }
private:
    sdrx_b200_detail::FirHandle f_;
};

#endif

// sdrx_b200 drop-in for sdrx's src/agc.hpp: classes AGC (agc.hpp:28-80) and LfAGC (agc.hpp:84-143) with the
// reference's constructors, setters, gain() and per-sample adjust(), backed by the CUDA library
// (sdrx_agc_adjust / sdrx_lfagc_adjust in ../sdrx_b200.h).  Include iqsample.hpp first, as the reference does.
//
// The recurrences run on the GPU in the reference's exact operation order.  The per-sample adjust() keeps
// the reference signature (one kernel launch per call: correct, slow); callers that own a whole chunk
// (sdrx.cpp:455-535 walks 512 samples per channel) should use the block overloads added here, or the
// batched engine (sdrx_engine_push_block) which runs every channel's AGC in one kernel.
#ifndef SDRX_B200_AGC_HPP
#define SDRX_B200_AGC_HPP

#include <cstdio>

#include <sdrx_b200.h>       // -I<repo>/include (the headers themselves may be copied next to sdrx.cpp)
#include "iqsample.hpp"

namespace sdrx_b200_detail {
inline void agc_init(sdrx_agc_state &s, float attack, float decay, float reference, float max_gain, float alpha,
                     float power) {
    s.attack = attack; s.decay = decay; s.reference = reference; s.max_gain = max_gain; s.alpha = alpha;
    s.min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;      // agc.hpp:34-35
    s.gain = 1.0f; s.power = power; s.active = 0;
}
inline void agc_check(int rc, const char *what) {
    if (rc != 0) std::fprintf(stderr, "sdrx_b200: %s failed (%d): %s\n", what, rc, sdrx_last_error());
}
}  // namespace sdrx_b200_detail

class AGC {
public:
    AGC(float attack = 10.0f, float decay = 0.01f, float reference = 0.25f, float max_gain = 200.0f, float alpha = 0.01f,
        int device = 0) : device_(device) {
        sdrx_b200_detail::agc_init(s_, attack, decay, reference, max_gain, alpha, reference);   // power_(reference_)
    }
    void setAttack(float attack) { s_.attack = attack; }
    void setDecay(float decay) { s_.decay = decay; }
    void setReference(float reference) { s_.reference = reference; }     // power_ keeps its value (agc.hpp:40)
    void setMaxGain(float max_gain) {
        s_.max_gain = max_gain;
        s_.min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
    }
    void setAlpha(float alpha) { s_.alpha = alpha; }
    float gain() { return s_.gain; }

    iqsample_t adjust(iqsample_t sample) {
        iqsample_t out(0.0f, 0.0f);
        adjust(&sample, 1, &out);
        return out;
    }
    // block form: n samples in order, in == out allowed
    void adjust(const iqsample_t *in, unsigned n, iqsample_t *out) {
        sdrx_b200_detail::agc_check(sdrx_agc_adjust(device_, &s_, reinterpret_cast<const float *>(in), n,
                                                    reinterpret_cast<float *>(out)), "sdrx_agc_adjust");
    }

private:
    sdrx_agc_state s_;
    int            device_;
};

class LfAGC {
public:
    LfAGC(float attack = 10.0f, float decay = 0.01f, float reference = 0.25f, float max_gain = 200.0f,
          float alpha = 0.005f, int device = 0) : device_(device) {
        sdrx_b200_detail::agc_init(s_, attack, decay, reference, max_gain, alpha, reference * reference);
    }
    void setAttack(float attack) { s_.attack = attack; }
    void setDecay(float decay) { s_.decay = decay; }
    void setReference(float reference) { s_.reference = reference; }
    void setMaxGain(float max_gain) {
        s_.max_gain = max_gain;
        s_.min_gain = (max_gain > 0.0f) ? 1.0f / max_gain : 1.0f / 200.0f;
    }
    void setAlpha(float alpha) { s_.alpha = alpha; }
    float gain() { return s_.gain; }
    bool active() { return s_.active != 0; }
    void activate() { s_.active = 1; }
    void deactivate() { s_.active = 0; }

    float adjust(float sample) {
        float out = 0.0f;
        adjust(&sample, 1, &out);
        return out;
    }
    void adjust(const float *in, unsigned n, float *out) {
        sdrx_b200_detail::agc_check(sdrx_lfagc_adjust(device_, &s_, in, n, out), "sdrx_lfagc_adjust");
    }

private:
    sdrx_agc_state s_;
    int            device_;
};

#endif

// sdrx_b200 drop-in for sdrx's src/msd.hpp: class MSD with the reference's public surface
// (msd.hpp:46-49 Stage, :52 MSD(), :54 MSD(translator, stages, use_ftfir), :71 m(), :75 decimate()),
// backed by the CUDA library through the C ABI (sdrx_msd_* in ../sdrx_b200.h).  Header only.
//
// Like the reference the class cannot fail at the call site (void returns, no exceptions): a CUDA error
// is reported once on stderr, remembered in status(), and decimate() then yields zero output samples.
// There is no CPU path.  MSD is copyable and a copy carries the delay lines and trans_pos_ of the
// original (sdrx.cpp:1745 copies Settings; ds.hpp:34 copies the MSD into its worker).
#ifndef SDRX_B200_MSD_HPP
#define SDRX_B200_MSD_HPP

#include <cstddef>
#include <cstdio>
#include <vector>

#include <sdrx_b200.h>       // -I<repo>/include (the headers themselves may be copied next to sdrx.cpp)
#include "iqsample.hpp"

class MSD {
public:
    struct Stage {
        size_t             m;    // down-sampling factor
        std::vector<float> h;    // low-pass FIR coefficients (odd, symmetric)
    };

    MSD() = default;

    MSD(const std::vector<iqsample_t> &translator, const std::vector<MSD::Stage> &stages, bool use_ftfir = false,
        int device = 0)
        : m_(1), translator_(translator), stages_(stages), use_ftfir_(use_ftfir), device_(device) {
        for (const Stage &s : stages_) m_ *= s.m;
        create_();
    }

    MSD(const MSD &o)
        : m_(o.m_), translator_(o.translator_), stages_(o.stages_), use_ftfir_(o.use_ftfir_), device_(o.device_),
          status_(o.status_), fed_(o.fed_) {
        if (o.h_) check_(sdrx_msd_clone(o.h_, &h_), "sdrx_msd_clone");
    }
    MSD(MSD &&o) noexcept { swap_(o); }
    MSD &operator=(MSD o) noexcept { swap_(o); return *this; }
    ~MSD() { sdrx_msd_destroy(h_); }

    // decimation factor of the whole chain
    size_t m() const { return m_; }

    // Translate and down-sample any number of samples; state is carried between calls.
    void decimate(const iqsample_t *in, size_t in_len, iqsample_t *out, size_t *out_len_ptr = nullptr) {
        size_t n = 0;
        if (h_ && in_len) {
            check_(sdrx_msd_decimate(h_, reinterpret_cast<const float *>(in), in_len, reinterpret_cast<float *>(out), &n),
                   "sdrx_msd_decimate");
            fed_ += in_len;
        }
        if (out_len_ptr) *out_len_ptr = n;
    }

    // ---- additions over the reference (used by ds.hpp to batch all channels on one engine)
    int status() const { return status_; }                      // 0 or the last negative sdrx_status
    int device() const { return device_; }
    bool fresh() const { return fed_ == 0; }                    // no sample seen yet
    bool use_ftfir() const { return use_ftfir_; }
    const std::vector<iqsample_t> &translator() const { return translator_; }
    const std::vector<Stage> &stages() const { return stages_; }

private:
    size_t                  m_ = 1;
    std::vector<iqsample_t> translator_;
    std::vector<Stage>      stages_;
    bool                    use_ftfir_ = false;
    int                     device_ = 0;
    int                     status_ = 0;
    size_t                  fed_ = 0;
    sdrx_msd               *h_ = nullptr;

    void create_() {
        if (stages_.empty()) return;
        std::vector<int> m, nt;
        std::vector<float> taps;
        for (const Stage &s : stages_) {
            m.push_back((int)s.m);
            nt.push_back((int)s.h.size());
            taps.insert(taps.end(), s.h.begin(), s.h.end());
        }
        check_(sdrx_msd_create(translator_.empty() ? nullptr : reinterpret_cast<const float *>(translator_.data()),
                               (int)translator_.size(), (int)stages_.size(), m.data(), nt.data(), taps.data(),
                               use_ftfir_ ? 1 : 0, device_, &h_),
               "sdrx_msd_create");
    }
    void check_(int rc, const char *what) {
        if (rc == 0) return;
        if (status_ == 0) std::fprintf(stderr, "sdrx_b200: %s failed (%d): %s\n", what, rc, sdrx_last_error());
        status_ = rc;
    }
    void swap_(MSD &o) noexcept {
        std::swap(m_, o.m_); translator_.swap(o.translator_); stages_.swap(o.stages_);
        std::swap(use_ftfir_, o.use_ftfir_); std::swap(device_, o.device_); std::swap(status_, o.status_);
        std::swap(fed_, o.fed_); std::swap(h_, o.h_);
    }
};

#endif

// sdrx_b200 drop-in for sdrx's src/ds.hpp: class DS (ds.hpp:32-78) -- DS(const MSD&), ~DS(),
// addJob(data, len, out, latch) -- re-thought for the GPU.  The reference gives every channel a worker
// thread that owns a copy of the channel's MSD; sdrx.cpp:327-340 hands the same IQ block to every DS and
// waits on one std::latch.  Here every live DS is one channel of a shared, batched CUDA engine
// (sdrx_engine_* in ../sdrx_b200.h): addJob() only records the job, and the call that completes the set
// (one job per live DS, all on the same block) pushes the block through the channel bank ONCE for all
// channels, scatters the per-channel output rows and counts every latch down.  No host threads are created.
//
// Contract (what sdrx.cpp already does): every live DS gets exactly one job per block and all jobs of a
// block name the same input.  Jobs that break it, MSDs that are not batchable (different stage lists or
// translator lengths, --ftfir) fall back to one sdrx_msd_decimate per channel, still on the GPU.
// Creating or destroying a DS after the stream has started restarts the bank's delay lines.
#ifndef SDRX_B200_DS_HPP
#define SDRX_B200_DS_HPP

#include <algorithm>
#include <cstring>
#include <latch>
#include <memory>
#include <mutex>
#include <vector>

#include "msd.hpp"

namespace sdrx_b200_detail {

class DsBank {
public:
    static DsBank &instance() { static DsBank b; return b; }

    int add(const MSD &msd) {
        std::lock_guard<std::mutex> g(mu_);
        slots_.push_back(Slot{true, msd, Job{}});
        ++n_live_;
        dirty_ = true;
        return (int)slots_.size() - 1;
    }
    void remove(int id) {
        std::lock_guard<std::mutex> g(mu_);
        Slot &s = slots_[id];
        if (!s.live) return;
        s.live = false;
        --n_live_;
        // the reference's ~DS joins its worker, which finishes a job already handed over (ds.hpp:36-41): run it, so a
        // thread blocked on that block's latch (sdrx.cpp:340) is released
        if (s.job.latch) { run_one_(s); --n_pending_; }
        dirty_ = true;
        if (n_live_ == 0) { sdrx_engine_destroy(eng_); eng_ = nullptr; slots_.clear(); }
        else if (n_pending_ == n_live_) run_all_();
    }
    void submit(int id, const iqsample_t *data, unsigned len, iqsample_t *out, std::shared_ptr<std::latch> latch) {
        std::lock_guard<std::mutex> g(mu_);
        Slot &s = slots_[id];
        if (s.job.latch) { run_one_(s); --n_pending_; }      // a second job before the set completed: keep order
        s.job = Job{data, len, out, std::move(latch)};
        if (++n_pending_ == n_live_) run_all_();
    }
    bool batched() { std::lock_guard<std::mutex> g(mu_); return eng_ != nullptr && !dirty_; }

private:
    struct Job { const iqsample_t *data = nullptr; unsigned len = 0; iqsample_t *out = nullptr; std::shared_ptr<std::latch> latch; };
    struct Slot { bool live; MSD msd; Job job; };
    std::mutex mu_;
    std::vector<Slot> slots_;
    std::vector<int> order_;          // live slot ids in engine channel order
    int n_live_ = 0, n_pending_ = 0;
    bool dirty_ = true, batchable_ = false;
    sdrx_engine *eng_ = nullptr;
    unsigned cap_ = 0;
    std::vector<float> rows_;

    static bool same_chain_(const MSD &a, const MSD &b) {
        if (a.stages().size() != b.stages().size() || a.device() != b.device()) return false;
        for (size_t i = 0; i < a.stages().size(); ++i)
            if (a.stages()[i].m != b.stages()[i].m || a.stages()[i].h != b.stages()[i].h) return false;
        return true;
    }
    void run_one_(Slot &s) {
        s.msd.decimate(s.job.data, s.job.len, s.job.out);
        s.job.latch->count_down();
        s.job = Job{};
    }
    // A block longer than the engine's capacity: a larger engine takes over WITH the delay lines and trans_pos of the
    // old one (MSD::decimate takes any in_len and keeps its state, msd.hpp:75) -- same as sdrx_msd_decimate's growth.
    bool grow_(unsigned len) {
        sdrx_engine *old = eng_;
        eng_ = nullptr;
        std::vector<unsigned char> st(old ? sdrx_engine_state_size(old) : 0);
        int rc = old ? sdrx_engine_save_state(old, st.data(), st.size()) : 0;
        const bool ok = rebuild_(len, false) && eng_ && !rc && old &&
                        sdrx_engine_load_state(eng_, st.data(), st.size()) == 0;
        sdrx_engine_destroy(old);
        if (!ok && eng_) {
            std::fprintf(stderr, "sdrx_b200: DS bank: could not carry the delay lines into the larger engine: %s\n", sdrx_last_error());
            batchable_ = false;
        }
        return ok;
    }
    bool rebuild_(unsigned len, bool destroy_old = true) {
        if (destroy_old) sdrx_engine_destroy(eng_);
        eng_ = nullptr;
        order_.clear();
        for (size_t i = 0; i < slots_.size(); ++i) if (slots_[i].live) order_.push_back((int)i);
        const MSD &first = slots_[order_[0]].msd;
        size_t N = 0;
        batchable_ = !first.stages().empty();
        for (int id : order_) {
            const MSD &m = slots_[id].msd;
            if (!same_chain_(first, m) || m.use_ftfir() || !m.fresh() || m.status() != 0) batchable_ = false;
            if (!m.translator().empty()) { if (N && N != m.translator().size()) batchable_ = false; N = m.translator().size(); }
        }
        dirty_ = false;
        if (!batchable_) return false;
        const int C = (int)order_.size();
        std::vector<float> tr((size_t)C * std::max<size_t>(N, 1) * 2, 0.0f);
        std::vector<uint8_t> has(C, 0);
        for (int c = 0; c < C; ++c) {
            const std::vector<iqsample_t> &t = slots_[order_[c]].msd.translator();
            if (t.empty()) continue;
            has[c] = 1;
            std::memcpy(tr.data() + (size_t)c * N * 2, t.data(), N * sizeof(iqsample_t));
        }
        std::vector<int> m, nt;
        std::vector<float> taps;
        for (const MSD::Stage &s : first.stages()) {
            m.push_back((int)s.m); nt.push_back((int)s.h.size());
            taps.insert(taps.end(), s.h.begin(), s.h.end());
        }
        sdrx_engine_config cfg;
        std::memset(&cfg, 0, sizeof(cfg));
        cfg.device = first.device(); cfg.n_channels = C; cfg.translator_len = (int)N;
        cfg.translators = N ? tr.data() : nullptr; cfg.has_translator = has.data();
        cfg.n_stages = (int)m.size(); cfg.m = m.data(); cfg.ntaps = nt.data(); cfg.taps = taps.data();
        cap_ = std::max(len, 1u << 16);
        cfg.block_len = cap_;
        cfg.flags = SDRX_FLAG_NO_DEMOD;
        if (sdrx_engine_create(&cfg, &eng_) != 0) {
            std::fprintf(stderr, "sdrx_b200: DS bank engine: %s\n", sdrx_last_error());
            eng_ = nullptr; batchable_ = false;
        }
        return batchable_;
    }
    void run_all_() {
        n_pending_ = 0;
        unsigned len = 0;
        const iqsample_t *data = nullptr;
        bool uniform = true;
        for (Slot &s : slots_) {
            if (!s.live) continue;
            if (!data) { data = s.job.data; len = s.job.len; }
            else if (s.job.data != data || s.job.len != len) uniform = false;
        }
        if (dirty_) rebuild_(len);
        else if (batchable_ && len > cap_) grow_(len);
        bool done = false;
        if (uniform && batchable_ && eng_ && len > 0) {
            unsigned n = 0;
            int rc = sdrx_engine_push_block(eng_, data, SDRX_FMT_F32, len);
            if (!rc) rc = sdrx_engine_last_out_len(eng_, &n);
            if (!rc && n) {
                rows_.resize((size_t)order_.size() * n * 2);
                rc = sdrx_engine_get_channel_iq(eng_, rows_.data());
            } else if (!rc) rc = sdrx_engine_sync(eng_);
            if (!rc) {
                for (size_t c = 0; c < order_.size(); ++c) {
                    Slot &s = slots_[order_[c]];
                    if (n) std::memcpy(static_cast<void *>(s.job.out), rows_.data() + c * n * 2, (size_t)n * sizeof(iqsample_t));
                    s.job.latch->count_down();
                    s.job = Job{};
                }
                done = true;
            } else {
                std::fprintf(stderr, "sdrx_b200: DS bank: %s\n", sdrx_last_error());
                batchable_ = false;
            }
        }
        // not batchable (or the batched pass failed): one sdrx_msd_decimate per channel.  After a FAILED batched pass the
        // per-slot MSD objects have not seen the stream so far, so their delay lines restart -- a CUDA failure is not
        // recoverable mid-stream, the message above says so; the latches are still counted down so nobody hangs.
        if (!done)
            for (Slot &s : slots_) if (s.live && s.job.latch) run_one_(s);
    }
};

}  // namespace sdrx_b200_detail

class DS {
public:
    DS(const MSD &msd) : id_(sdrx_b200_detail::DsBank::instance().add(msd)) {}
    DS(const DS &) = delete;
    DS &operator=(const DS &) = delete;
    ~DS() { sdrx_b200_detail::DsBank::instance().remove(id_); }

    void addJob(const iqsample_t *data, unsigned data_len, iqsample_t *out, std::shared_ptr<std::latch> latch) {
        sdrx_b200_detail::DsBank::instance().submit(id_, data, data_len, out, std::move(latch));
    }

private:
    int id_;
};

#endif

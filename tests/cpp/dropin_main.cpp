// Test driver for the C++ drop-in headers (include/sdrx_b200/*.hpp): uses ONLY the reference's class
// surface (MSD, DS, AGC, LfAGC, FIR2, FIR3<iqsample_t>) the way sdrx.cpp does, on data a Python test
// wrote, and dumps the results for comparison with the oracle (tests/test_gpu_dropin_cpp.py).
//   usage: dropin_test <spec.bin> <out.bin>
// spec: int32 C, N, S, block_len, n_blocks; int32 m[S]; int32 ntaps[S]; float taps[sum]; per channel:
//       int32 has_translator, float translator[N][2]; int32 n_fir2; float fir2[n_fir2];
//       float iq[n_blocks][block_len][2]
// out:  float msd[n_blocks][C][512][2]; float ds[n_blocks][C][512][2]; float agc[n_blocks][512][2];
//       float lf[n_blocks][512]; float fir2[n_blocks][1024]; float fir3[n_blocks][512][2]; int32 batched
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <latch>
#include <memory>
#include <vector>

#include <sdrx_b200/iqsample.hpp>
#include <sdrx_b200/msd.hpp>
#include <sdrx_b200/ds.hpp>
#include <sdrx_b200/agc.hpp>
#include <sdrx_b200/fir.hpp>

template <typename T>
static std::vector<T> rd(std::ifstream &f, size_t n) {
    std::vector<T> v(n);
    f.read(reinterpret_cast<char *>(v.data()), (std::streamsize)(n * sizeof(T)));
    if (!f) { std::cerr << "short spec file\n"; std::exit(2); }
    return v;
}
template <typename T>
static void wr(std::ofstream &f, const T *p, size_t n) { f.write(reinterpret_cast<const char *>(p), (std::streamsize)(n * sizeof(T))); }

int main(int argc, char **argv) {
    if (argc != 3) { std::cerr << "usage: dropin_test spec out\n"; return 2; }
    std::ifstream f(argv[1], std::ios::binary);
    auto hdr = rd<int>(f, 5);
    const int C = hdr[0], N = hdr[1], S = hdr[2], L = hdr[3], NB = hdr[4];
    auto m = rd<int>(f, S), nt = rd<int>(f, S);
    std::vector<MSD::Stage> stages;
    for (int s = 0; s < S; ++s) stages.push_back(MSD::Stage{(size_t)m[s], rd<float>(f, nt[s])});
    std::vector<MSD> msds;                       // sdrx.cpp:1692  ch.msd = MSD(translator, stages, use_ftfir)
    for (int c = 0; c < C; ++c) {
        const int has = rd<int>(f, 1)[0];
        auto t = rd<iqsample_t>(f, N);
        msds.push_back(MSD(has ? t : std::vector<iqsample_t>(), stages));
    }
    const int nf = rd<int>(f, 1)[0];
    auto fir2c = rd<float>(f, nf);
    auto iq = rd<iqsample_t>(f, (size_t)NB * L);

    std::vector<std::unique_ptr<DS>> ds;         // sdrx.cpp:1694-1698  new DS(ch.msd)
    for (int c = 0; c < C; ++c) ds.push_back(std::make_unique<DS>(msds[c]));
    std::vector<MSD> seq = msds;                 // sdrx.cpp:1745: Settings (and every MSD) is copied

    AGC agc;                                     // sdrx.cpp:1702-1705
    agc.setReference(1.0f); agc.setAttack(1.0f); agc.setDecay(0.01f); agc.setMaxGain(300.0f);
    LfAGC lf;                                    // sdrx.cpp:1707-1710
    lf.setReference(1.0f); lf.setAttack(1.0f); lf.setDecay(0.01f); lf.activate();
    FIR2 fir2(fir2c);                            // sdrx.cpp:880-881
    fir2.setGain(3.0f);
    FIR3<iqsample_t> fir3(fir2c);

    std::ofstream o(argv[2], std::ios::binary);
    std::vector<iqsample_t> a((size_t)NB * C * 512), b((size_t)NB * C * 512), g((size_t)NB * 512), f3((size_t)NB * 512);
    std::vector<float> lfo((size_t)NB * 512), f2((size_t)NB * 1024);
    for (int blk = 0; blk < NB; ++blk) {
        const iqsample_t *data = iq.data() + (size_t)blk * L;
        iqsample_t *row = a.data() + (size_t)blk * C * 512;
        for (int c = 0; c < C; ++c) seq[c].decimate(data, L, row + (size_t)c * 512);     // sdrx.cpp:342-345
        iqsample_t *row2 = b.data() + (size_t)blk * C * 512;
        auto latch = std::make_shared<std::latch>(C);                                     // sdrx.cpp:327-340
        for (int c = 0; c < C; ++c) ds[c]->addJob(data, L, row2 + (size_t)c * 512, latch);
        latch->wait();
        iqsample_t *gp = g.data() + (size_t)blk * 512;
        gp[0] = agc.adjust(row[0]);                                                       // per-sample form, sdrx.cpp:457
        agc.adjust(row + 1, 511, gp + 1);                                                 // block form
        float env[512];                                   // |re|+|im| is exact in FP32: the Python side restates it
        for (int i = 0; i < 512; ++i) env[i] = std::fabs(gp[i].real()) + std::fabs(gp[i].imag());
        float *lp = lfo.data() + (size_t)blk * 512;
        lp[0] = lf.adjust(env[0]);
        lf.adjust(env + 1, 511, lp + 1);
        float *st = f2.data() + (size_t)blk * 1024;
        for (int i = 0; i < 512; ++i) { st[2 * i] = 0.8f * lp[i]; st[2 * i + 1] = 0.2f * lp[i]; }
        fir2.filter(st, 1024, st);                                                        // in place, sdrx.cpp:646
        fir3.filter(row, 512, f3.data() + (size_t)blk * 512);
    }
    const int batched = sdrx_b200_detail::DsBank::instance().batched() ? 1 : 0;
    wr(o, a.data(), a.size()); wr(o, b.data(), b.size()); wr(o, g.data(), g.size());
    wr(o, lfo.data(), lfo.size()); wr(o, f2.data(), f2.size()); wr(o, f3.data(), f3.size());
    wr(o, &batched, 1);
    int bad = 0;
    for (auto &q : seq) bad |= q.status();
    std::printf("dropin_test: %d channels x %d blocks, DS bank batched=%d, status=%d\n", C, NB, batched, bad);
    return bad ? 1 : 0;
}

// sdrx_b200 -- single-object C ABI behind the drop-in C++ classes (include/sdrx_b200/msd.hpp, ds.hpp, agc.hpp, fir.hpp).
// An MSD object is a one-channel channelizer engine; AGC/LfAGC/FIR objects run their exact
// sequential kernels on caller-supplied arrays.  Same numerics as the batched engine.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "host_common.h"
#include "demod.cuh"

using namespace sdrx;

// ---------------------------------------------------------------------------------- MSD (msd.hpp)
struct sdrx_msd {
    int device = 0;
    int N = 0, S = 0;
    std::vector<float> translator;          // [N][2], empty when no tuning (msd.hpp:79)
    std::vector<int> m, ntaps;
    std::vector<float> taps;
    size_t m_total = 1;
    bool use_ftfir = false;                 // msd.hpp:54 third argument
    unsigned cap = 0;                       // block_len of the current engine
    sdrx_engine *eng = nullptr;
};

namespace {

int msd_make_engine(sdrx_msd *q, unsigned cap, sdrx_engine **out) {
    sdrx_engine_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.device = q->device;
    cfg.n_channels = 1;
    cfg.translator_len = q->N;
    uint8_t has = q->translator.empty() ? 0 : 1;
    cfg.translators = q->translator.empty() ? nullptr : q->translator.data();
    cfg.has_translator = &has;
    cfg.n_stages = q->S;
    cfg.m = q->m.data();
    cfg.ntaps = q->ntaps.data();
    cfg.taps = q->taps.data();
    cfg.block_len = cap;
    cfg.flags = SDRX_FLAG_NO_DEMOD | SDRX_FLAG_FORCE_GENERIC | (q->use_ftfir ? SDRX_FLAG_FTFIR : 0u);
    return sdrx_engine_create(&cfg, out);
}

// grow the engine's block capacity, carrying the delay lines over (MSD::decimate takes any in_len)
int msd_reserve(sdrx_msd *q, size_t in_len) {
    if (q->eng && in_len <= q->cap) return SDRX_OK;
    unsigned cap = (unsigned)std::max<size_t>(in_len, std::max<size_t>(q->cap * 2, 1u << 16));
    sdrx_engine *ne = nullptr;
    int rc = msd_make_engine(q, cap, &ne);
    if (rc) return rc;
    if (q->eng) {
        std::vector<uint8_t> st(sdrx_engine_state_size(q->eng));
        rc = sdrx_engine_save_state(q->eng, st.data(), st.size());
        if (!rc) rc = sdrx_engine_load_state(ne, st.data(), st.size());
        if (rc) { sdrx_engine_destroy(ne); return rc; }
        sdrx_engine_destroy(q->eng);
    }
    q->eng = ne;
    q->cap = cap;
    return SDRX_OK;
}

}  // namespace

extern "C" int sdrx_msd_create(const float *translator, int N, int n_stages, const int *m, const int *ntaps,
                               const float *taps, int use_ftfir, int device, sdrx_msd **out) {
    if (!out || !m || !ntaps || !taps || n_stages < 1) return fail(SDRX_ERR_ARG, "null or empty stage list");
    *out = nullptr;
    sdrx_msd *q = new sdrx_msd();
    q->device = device;
    q->S = n_stages;
    q->use_ftfir = use_ftfir != 0;
    q->N = (translator && N > 0) ? N : 0;
    if (q->N) q->translator.assign(translator, translator + 2 * (size_t)N);
    int tot = 0;
    for (int s = 0; s < n_stages; ++s) {
        q->m.push_back(m[s]);
        q->ntaps.push_back(ntaps[s]);
        q->m_total *= (size_t)m[s];
        tot += ntaps[s];
    }
    q->taps.assign(taps, taps + tot);
    int rc = msd_reserve(q, 1u << 16);
    if (rc) { delete q; return rc; }
    *out = q;
    return SDRX_OK;
}

extern "C" int sdrx_msd_clone(const sdrx_msd *src, sdrx_msd **out) {
    if (!src || !out) return fail(SDRX_ERR_ARG, "null argument");
    *out = nullptr;
    sdrx_msd *q = new sdrx_msd();
    q->device = src->device; q->N = src->N; q->S = src->S;
    q->translator = src->translator; q->m = src->m; q->ntaps = src->ntaps; q->taps = src->taps;
    q->m_total = src->m_total;
    q->use_ftfir = src->use_ftfir;
    int rc = msd_make_engine(q, src->cap, &q->eng);
    if (rc) { delete q; return rc; }
    q->cap = src->cap;
    std::vector<uint8_t> st(sdrx_engine_state_size(src->eng));
    rc = sdrx_engine_save_state(src->eng, st.data(), st.size());
    if (!rc) rc = sdrx_engine_load_state(q->eng, st.data(), st.size());
    if (rc) { sdrx_engine_destroy(q->eng); delete q; return rc; }
    *out = q;
    return SDRX_OK;
}

extern "C" void sdrx_msd_destroy(sdrx_msd *q) {
    if (!q) return;
    sdrx_engine_destroy(q->eng);
    delete q;
}

extern "C" size_t sdrx_msd_m(const sdrx_msd *q) { return q ? q->m_total : 0; }

extern "C" int sdrx_msd_decimate(sdrx_msd *q, const float *in, size_t in_len, float *out, size_t *out_len) {
    if (!q || (!in && in_len) || !out) return fail(SDRX_ERR_ARG, "null argument");
    if (out_len) *out_len = 0;
    if (in_len == 0) return SDRX_OK;
    int rc = msd_reserve(q, in_len);
    if (rc) return rc;
    rc = sdrx_engine_push_block(q->eng, in, SDRX_FMT_F32, (unsigned)in_len);
    if (rc) return rc;
    unsigned n = 0;
    sdrx_engine_last_out_len(q->eng, &n);
    rc = n ? sdrx_engine_get_channel_iq(q->eng, out) : sdrx_engine_sync(q->eng);
    if (rc) return rc;
    if (out_len) *out_len = n;
    return SDRX_OK;
}

// ------------------------------------------------------------------------- AGC / LfAGC (agc.hpp)
namespace {

__global__ void k_agc_adjust(const float2 *__restrict__ in, unsigned n, float2 *__restrict__ out, AgcParams a,
                             float *__restrict__ gp /* gain, power */) {
    float g = gp[0], p = gp[1];
    for (unsigned i = 0; i < n; ++i) {                              // agc.hpp:48-69
        const float2 x = in[i];
        const float2 y = make_float2(__fmul_rn(x.x, g), __fmul_rn(x.y, g));
        p = __fadd_rn(p, __fmul_rn(a.alpha, __fsub_rn(norm_exact(y), p)));
        const float err = __fsub_rn(a.reference, p);
        g = __fadd_rn(g, __fmul_rn(err > 0.0f ? a.decay : a.attack, err));
        g = clamp_gain(g, a.min_gain, a.max_gain);
        out[i] = y;
    }
    gp[0] = g; gp[1] = p;
}

__global__ void k_lfagc_adjust(const float *__restrict__ in, unsigned n, float *__restrict__ out, AgcParams a,
                               int active, float *__restrict__ gp) {
    float g = gp[0], p = gp[1];
    for (unsigned i = 0; i < n; ++i) {                              // agc.hpp:106-131
        const float s = in[i];
        const float y = __fmul_rn(s, g);
        p = __fadd_rn(p, __fmul_rn(a.alpha, __fsub_rn(__fmul_rn(y, y), p)));
        const float err = __fsub_rn(a.reference, __fsqrt_rn(p));
        g = __fadd_rn(g, __fmul_rn(err > 0.0f ? a.decay : a.attack, err));
        g = clamp_gain(g, a.min_gain, a.max_gain);
        out[i] = active ? y : s;
    }
    gp[0] = g; gp[1] = p;
}

// Scratch of the AGC entry points: the state struct is the caller's plain data, so the device buffers live here, one
// grow-only set per device, behind a mutex (AGC::adjust is called per sample by sdrx.cpp:457 -- no allocation per call).
struct AgcScratch { float *d_in = nullptr, *d_out = nullptr, *d_gp = nullptr; size_t cap = 0; };
std::mutex g_agc_mu;
AgcScratch g_agc_scratch[64];

template <bool LF>
int agc_run(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out) {
    if (!s || (!in && n) || (!out && n)) return fail(SDRX_ERR_ARG, "null argument");
    if (n == 0) return SDRX_OK;
    if (device < 0 || device >= 64) return fail(SDRX_ERR_ARG, "device ordinal out of range");
    DevGuard g(device);
    if (!g.ok) return fail(SDRX_ERR_CUDA, "no CUDA device: sdrx_b200 has no CPU path");
    const size_t elems = LF ? n : 2 * (size_t)n;
    std::lock_guard<std::mutex> lk(g_agc_mu);
    AgcScratch &sc = g_agc_scratch[device];
    if (!sc.d_gp) CU(cudaMalloc(&sc.d_gp, 8));
    if (elems > sc.cap) {
        cudaFree(sc.d_in); cudaFree(sc.d_out);
        sc.d_in = sc.d_out = nullptr; sc.cap = 0;
        const size_t cap = std::max<size_t>(elems, 4096);
        CU(cudaMalloc(&sc.d_in, cap * 4));
        CU(cudaMalloc(&sc.d_out, cap * 4));
        sc.cap = cap;
    }
    float gp[2] = {s->gain, s->power};
    CU(cudaMemcpy(sc.d_in, in, elems * 4, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(sc.d_gp, gp, 8, cudaMemcpyHostToDevice));
    AgcParams a{s->attack, s->decay, s->reference, s->min_gain, s->max_gain, s->alpha};
    if (LF) k_lfagc_adjust<<<1, 1>>>(sc.d_in, n, sc.d_out, a, s->active, sc.d_gp);
    else    k_agc_adjust<<<1, 1>>>((const float2 *)sc.d_in, n, (float2 *)sc.d_out, a, sc.d_gp);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, sc.d_out, elems * 4, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(gp, sc.d_gp, 8, cudaMemcpyDeviceToHost));
    s->gain = gp[0]; s->power = gp[1];
    return SDRX_OK;
}

}  // namespace

extern "C" int sdrx_agc_adjust(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out) {
    return agc_run<false>(device, s, in, n, out);
}
extern "C" int sdrx_lfagc_adjust(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out) {
    return agc_run<true>(device, s, in, n, out);
}

// ---- sdrx_agc_adjust without any allocation (no counterpart in the reference)
// The scratch sdrx_agc_adjust keeps in grow-only cudaMalloc'ed buffers -- the staged input, the staged output and the
// (gain, power) pair -- lives in ONE file-scope device array of a fixed size instead: every GPU that loads the library
// holds these 129 KB until exit.  In host code `g_agc_static` is only a stand-in for the symbol; the device address comes
// from cudaGetSymbolAddress, once per device, with that device current.
constexpr unsigned kAgcStaticMax = SDRX_AGC_NOALLOC_MAX;          // complex samples per call
static __device__ __align__(256) float g_agc_static[2 * kAgcStaticMax /* in */ + 2 * kAgcStaticMax /* out */ + 64 /* gain, power */];
namespace {
std::mutex g_agc_static_mu;
float *g_agc_static_ptr[64] = {};
}  // namespace

extern "C" int sdrx_agc_adjust_noalloc(int device, sdrx_agc_state *s, const float *in, unsigned n, float *out) {
    if (!s || (!in && n) || (!out && n)) return fail(SDRX_ERR_ARG, "null argument");
    if (n > kAgcStaticMax)
        return fail(SDRX_ERR_ARG, "sdrx_agc_adjust_noalloc: " + std::to_string(n) + " samples in one call, the static scratch holds " +
                                      std::to_string(kAgcStaticMax) + " (SDRX_AGC_NOALLOC_MAX): split the call or use sdrx_agc_adjust");
    if (n == 0) return SDRX_OK;
    if (device < 0 || device >= 64) return fail(SDRX_ERR_ARG, "device ordinal out of range");
    DevGuard g(device);
    if (!g.ok) return fail(SDRX_ERR_CUDA, "no CUDA device: sdrx_b200 has no CPU path");
    std::lock_guard<std::mutex> lk(g_agc_static_mu);             // calls share the one array: serialised here, and synchronous below
    float *base = g_agc_static_ptr[device];
    if (!base) {
        CU(cudaGetSymbolAddress((void **)&base, g_agc_static));  // this device's copy of the array (the device is current)
        g_agc_static_ptr[device] = base;
    }
    float *d_in = base, *d_out = base + 2 * kAgcStaticMax, *d_gp = base + 4 * kAgcStaticMax;
    // nothing has to start at zero: the n input samples and (gain, power) are overwritten here, the kernel writes all n outputs,
    // and only those are read back -- what an earlier call left behind in the rest of the array is never looked at
    float gp[2] = {s->gain, s->power};
    CU(cudaMemcpy(d_in, in, (size_t)n * 8, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_gp, gp, 8, cudaMemcpyHostToDevice));
    AgcParams a{s->attack, s->decay, s->reference, s->min_gain, s->max_gain, s->alpha};
    k_agc_adjust<<<1, 1>>>((const float2 *)d_in, n, (float2 *)d_out, a, d_gp);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(gp, d_gp, 8, cudaMemcpyDeviceToHost));
    s->gain = gp[0]; s->power = gp[1];
    return SDRX_OK;
}

// ---------------------------------------------------------------------- FIR / FIR2 / FIR3 (fir.hpp)
struct sdrx_fir {
    int kind = 1, device = 0, n = 0;
    int lanes = 1;                       // interleaved streams sharing the taps: FIR 1, FIR2 2, FIR3<iq> 2
    std::vector<float> c, c_adj;
    float gain_db = 0.0f;
    float *d_c = nullptr, *d_hist = nullptr;     // hist: (n-1) frames of `lanes` floats, oldest first
    float *d_buf = nullptr, *d_out = nullptr;    // grow-only scratch of filter(): [hist | new] frames, outputs
    size_t buf_cap = 0, out_cap = 0;
};

namespace {

// out[f][l] = sum_i c_adj[i] * frame[f + i][l] over [hist | new]; c_adj[0] meets the oldest sample
// (fir.hpp:40-44 after the ring pointer has advanced).  Exact, sequential in i.
__global__ void __launch_bounds__(256)
k_fir_frames(const float *__restrict__ buf /* [(hist+frames)][lanes] */, int lanes, int frames, const float *__restrict__ c,
             int ntaps, float *__restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= frames * lanes) return;
    const float *b = buf + t;
    float o = 0.0f;
    for (int k = 0; k < ntaps; ++k) o = __fadd_rn(o, __fmul_rn(c[k], b[(long)k * lanes]));
    out[t] = o;
}

}  // namespace

extern "C" int sdrx_fir_create(int kind, const float *c, int n, int device, sdrx_fir **out) {
    if (!out || !c || n < 1 || kind < 1 || kind > 3) return fail(SDRX_ERR_ARG, "bad FIR arguments");
    *out = nullptr;
    DevGuard g(device);
    if (!g.ok) return fail(SDRX_ERR_CUDA, "no CUDA device: sdrx_b200 has no CPU path");
    sdrx_fir *f = new sdrx_fir();
    f->kind = kind; f->device = device; f->n = n; f->lanes = kind == 1 ? 1 : 2;
    f->c.assign(c, c + n);
    f->c_adj = f->c;
    cudaError_t e1 = cudaMalloc(&f->d_c, sizeof(float) * n);
    cudaError_t e2 = cudaMalloc(&f->d_hist, sizeof(float) * std::max(1, (n - 1) * f->lanes));
    if (e1 != cudaSuccess || e2 != cudaSuccess) { cudaFree(f->d_c); cudaFree(f->d_hist); delete f; return fail(SDRX_ERR_CUDA, "cudaMalloc failed"); }
    cudaMemcpy(f->d_c, f->c_adj.data(), sizeof(float) * n, cudaMemcpyHostToDevice);
    cudaMemset(f->d_hist, 0, sizeof(float) * std::max(1, (n - 1) * f->lanes));
    *out = f;
    return SDRX_OK;
}

extern "C" void sdrx_fir_destroy(sdrx_fir *f) {
    if (!f) return;
    DevGuard g(f->device);
    cudaFree(f->d_c); cudaFree(f->d_hist); cudaFree(f->d_buf); cudaFree(f->d_out);
    delete f;
}

// The reference's FIR classes are implicitly copyable and a copy carries the delay line (buf_ and pos_ are plain
// members, fir.hpp:69-74, :121-127, :176-182): coefficients, gain and the device-side history.
extern "C" int sdrx_fir_clone(const sdrx_fir *src, sdrx_fir **out) {
    if (!src || !out) return fail(SDRX_ERR_ARG, "null argument");
    *out = nullptr;
    sdrx_fir *f = nullptr;
    int rc = sdrx_fir_create(src->kind, src->c.data(), src->n, src->device, &f);
    if (rc) return rc;
    DevGuard g(src->device);
    f->gain_db = src->gain_db;
    f->c_adj = src->c_adj;
    const size_t hist = (size_t)std::max(1, (src->n - 1) * src->lanes);
    cudaError_t e1 = cudaMemcpy(f->d_c, f->c_adj.data(), sizeof(float) * f->n, cudaMemcpyHostToDevice);
    cudaError_t e2 = cudaMemcpy(f->d_hist, src->d_hist, sizeof(float) * hist, cudaMemcpyDeviceToDevice);
    if (e1 != cudaSuccess || e2 != cudaSuccess) { sdrx_fir_destroy(f); cudaGetLastError(); return fail(SDRX_ERR_CUDA, "sdrx_fir_clone: copy failed"); }
    *out = f;
    return SDRX_OK;
}

extern "C" int sdrx_fir_set_gain(sdrx_fir *f, float gain_db) {   // fir.hpp:51-60
    if (!f) return fail(SDRX_ERR_ARG, "null argument");
    DevGuard g(f->device);
    f->gain_db = gain_db;
    for (int i = 0; i < f->n; ++i) f->c_adj[i] = f->c[i] * std::pow(10.0f, f->gain_db / 20.0f);
    CU(cudaMemcpy(f->d_c, f->c_adj.data(), sizeof(float) * f->n, cudaMemcpyHostToDevice));
    return SDRX_OK;
}

// in_len follows the reference: floats for FIR and FIR2 (fir.hpp:85, :137), complex samples for FIR3<iq>.
extern "C" int sdrx_fir_filter(sdrx_fir *f, const float *in, unsigned in_len, float *out) {
    if (!f || (!in && in_len) || (!out && in_len)) return fail(SDRX_ERR_ARG, "null argument");
    if (in_len == 0) return SDRX_OK;
    DevGuard g(f->device);
    const int lanes = f->lanes;
    const int frames = f->kind == 3 ? (int)in_len : (int)in_len / lanes;
    const size_t hist = (size_t)(f->n - 1) * lanes, fresh = (size_t)frames * lanes;
    if (hist + fresh > f->buf_cap) {
        cudaFree(f->d_buf); f->d_buf = nullptr; f->buf_cap = 0;
        CU(cudaMalloc(&f->d_buf, sizeof(float) * (hist + fresh)));
        f->buf_cap = hist + fresh;
    }
    if (fresh > f->out_cap) {
        cudaFree(f->d_out); f->d_out = nullptr; f->out_cap = 0;
        CU(cudaMalloc(&f->d_out, sizeof(float) * fresh));
        f->out_cap = fresh;
    }
    float *d_buf = f->d_buf, *d_out = f->d_out;
    if (hist) CU(cudaMemcpy(d_buf, f->d_hist, sizeof(float) * hist, cudaMemcpyDeviceToDevice));
    CU(cudaMemcpy(d_buf + hist, in, sizeof(float) * fresh, cudaMemcpyHostToDevice));
    k_fir_frames<<<div_up((long)fresh, 256), 256>>>(d_buf, lanes, frames, f->d_c, f->n, d_out);
    CU(cudaGetLastError());
    CU(cudaMemcpy(out, d_out, sizeof(float) * fresh, cudaMemcpyDeviceToHost));
    if (hist) CU(cudaMemcpy(f->d_hist, d_buf + fresh, sizeof(float) * hist, cudaMemcpyDeviceToDevice));
    return SDRX_OK;
}

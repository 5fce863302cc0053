#!/usr/bin/env python
"""CPU emulation (numpy, no GPU): how close to the oracle can fold + table GEMM come on the channels it is worst on?
Variants of the arithmetic, all on the bench's config-3 blocks, a subset of channels (noise-only channels next to the
strongest carriers, where the FP32 forms reach 1.35e-4, plus carrier channels and a random sample):
  exact        fold and GEMM in float64 (the correctly rounded linear map of the reference's own float tables and taps)
  fold32       fold as the GPU does it today (FP32 fma chain, newest tap first), GEMM in float64
  fold32_u16   ... and the GEMM as FP32 partial sums over 16 table entries added in FP32 (what k_tgemm / k_tgemm_tc do)
  exact_u16    float64 fold rounded to FP32 (= three bf16 planes), FP32 16-entry partial sums added in FP32
  exact_u16k   ... the partial sums added with compensation (float64 stands in for Kahan)
usage: python tools/emul_exact_gemm.py [--blocks 2] [--out file.json]"""
import argparse, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bindings as ob
from sdrx_b200 import config as cf, workloads as wl
from tests.harness import product_tables, rel_rms

ap = argparse.ArgumentParser()
ap.add_argument("--blocks", type=int, default=2)
ap.add_argument("--out", default=None)
ap.add_argument("--all", action="store_true", help="all 960 channels")
a = ap.parse_args()
plan, carriers, desc = wl.config3(960)
blocks = wl.synth_blocks(carriers, a.blocks, 1)
rc = cf.rate_config(10000000)
N, D = rc.N, rc.M
hc = np.ones(1)
up = 1
for m, taps in rc.stages:                                   # h_c(z) = prod h_s(z^(m_0 .. m_(s-1))), float taps, exact in float64
    t = np.zeros((len(taps) - 1) * up + 1)
    t[::up] = taps.astype(np.float64)
    hc = np.convolve(hc, t)
    up *= m
Lh = hc.size
# ---- channels: neighbours of the strongest carriers, the carriers themselves, a random sample
amp = {c.offset: c.amp for c in carriers}
strong = sorted(amp, key=lambda o: -amp[o])[:12]
offs = set()
for o in strong:
    for d in (-2, -1, 1, 2):
        if -480 <= o + d < 480 and (o + d) not in amp:
            offs.add(o + d)
offs |= set(strong[:6])
offs |= set(np.random.default_rng(3).choice(np.arange(-480, 480), 24, replace=False).tolist())
offs = sorted(offs) if not a.all else list(range(-480, 480))
print(f"{len(offs)} channels, {a.blocks} blocks; prototype {Lh} taps, D {D}, N {N}", flush=True)
orx = ob.OracleRx(product_tables(10000000), offs)
T = np.stack([cf.make_translator(N, rc.z, o) if o else np.ones(N, np.complex64) for o in offs])     # [C][N] the reference's float tables
T64 = T.astype(np.complex128)

hist = np.zeros(Lh - D, np.complex64)
pos0 = 0
res = {k: [] for k in ("exact", "fold32", "fold32_u16", "exact_u16", "exact_u16k")}
hc32 = hc.astype(np.float32)
for bi, blk in enumerate(blocks):
    want = orx.process(blk, 2).ch_iq                         # [C][512]
    x = np.concatenate([hist, blk])                          # x[h + n], n = sample index in the block
    h = hist.size
    n_out = blk.size // D
    U64 = np.zeros((N, n_out), np.complex128)
    U32 = np.zeros((N, n_out), np.complex64)
    for M in range(n_out):
        nM = D * M + D - 1
        w = x[h + nM - Lh + 1: h + nM + 1][::-1]             # w[j] = x[nM - j]
        p = (pos0 + nM - np.arange(Lh)) % N                  # table entry the reference multiplies that sample with
        prod = hc * w.astype(np.complex128)
        U64[:, M] = np.bincount(p, prod.real, N) + 1j * np.bincount(p, prod.imag, N)
        # the GPU's fold: per bin, taps j = d, d + N, ... (newest sample first), FP32 fma chain
        acc_r = np.zeros(N, np.float32); acc_i = np.zeros(N, np.float32)
        d = (pos0 + nM - np.arange(N)) % N                   # distance back from nM to the newest sample of bin p
        for q in range((Lh + N - 1) // N):
            j = d + q * N
            ok = j < Lh
            jj = np.where(ok, j, 0)
            hv = np.where(ok, hc32[jj], np.float32(0))
            xv = np.where(ok, w[jj], np.complex64(0))
            # fmaf(h, x, acc): one rounding -- float64 product of two floats is exact, the sum rounds once to FP32
            acc_r = (hv.astype(np.float64) * xv.real.astype(np.float64) + acc_r.astype(np.float64)).astype(np.float32)
            acc_i = (hv.astype(np.float64) * xv.imag.astype(np.float64) + acc_i.astype(np.float64)).astype(np.float32)
        U32[:, M] = acc_r + 1j * acc_i

    def units_fp32(Tm, U, comp):
        """FP32 partial sums over 16 entries (each accumulated in FP32 from exact products, term by term), then the 75 unit
        sums added in FP32 (comp False) or exactly (comp True)."""
        Tm = Tm.astype(np.complex64); U = U.astype(np.complex64)
        tot = np.zeros((Tm.shape[0], U.shape[1]), np.complex128 if comp else np.complex64)
        for u0 in range(0, N, 16):
            pr = np.zeros(tot.shape, np.float32); pi = np.zeros(tot.shape, np.float32)
            for k in range(u0, u0 + 16):
                t_, u_ = Tm[:, k:k + 1].astype(np.complex128), U[k:k + 1, :].astype(np.complex128)
                pr = (pr.astype(np.float64) + t_.real * u_.real).astype(np.float32)
                pr = (pr.astype(np.float64) - t_.imag * u_.imag).astype(np.float32)
                pi = (pi.astype(np.float64) + t_.real * u_.imag).astype(np.float32)
                pi = (pi.astype(np.float64) + t_.imag * u_.real).astype(np.float32)
            tot = tot + (pr + 1j * pi).astype(tot.dtype)
        return tot
    Y = {"exact": T64 @ U64, "fold32": T64 @ U32.astype(np.complex128), "fold32_u16": units_fp32(T, U32, False),
         "exact_u16": units_fp32(T, U64.astype(np.complex64), False), "exact_u16k": units_fp32(T, U64.astype(np.complex64), True)}
    for k, y in Y.items():
        res[k].append([rel_rms(y[c].astype(np.complex64), want[c]) for c in range(len(offs))])
    print(f"block {bi}: " + "  ".join(f"{k} max {max(res[k][-1]):.2e}" for k in res), flush=True)
    hist = x[-(Lh - D):]
    pos0 = (pos0 + blk.size) % N
out = {"workload": desc, "channels": offs, "blocks": a.blocks, "bar": 1e-4}
for k, v in res.items():
    v = np.asarray(v)
    out[k] = {"max": float(v.max()), "median": float(np.median(v)), "above_1e-4": int((v > 1e-4).sum()), "worst_channel_offset": int(offs[int(v.max(axis=0).argmax())])}
    print(k, out[k])
if a.out:
    json.dump(out, open(a.out, "w"), indent=1)

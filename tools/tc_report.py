#!/usr/bin/env python
"""Tensor-core table GEMM (SDRX_FLAG_TGEMM_TC, tgemm_tc.cuh) against the oracle, beside the FP32 table GEMM and the
time-domain bank: per-channel error, squelch-decision agreement, audio error, kernel time and whole-chain step, for the
plane-product / drain-unit variants (SDRX_TC_PRODUCTS, SDRX_TC_STEPS).   python tools/tc_report.py [--out file.json]"""
import argparse, ctypes as C, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import bindings as ob
from sdrx_b200 import Engine, _lib, workloads as wl
from tests.harness import product_tables, rel_rms

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=None)
ap.add_argument("--blocks", type=int, default=8)
ap.add_argument("--channels", type=int, default=960)
ap.add_argument("--threads", type=int, default=8)
ap.add_argument("--variants", default="9x1,9x2,9x4,6x1,6x4")
a = ap.parse_args()
plan, carriers, desc = wl.config3(a.channels)
blocks = wl.synth_blocks(carriers, a.blocks, 1)
orx = ob.OracleRx(product_tables(10000000), plan.offsets)
want = [orx.process(b, 2) for b in blocks]
car = sorted({c.offset for c in carriers})
is_car = np.asarray([o in car for o in plan.offsets])
ever_open = np.stack([w.sql for w in want]).any(axis=0)

engines = [("bank", lambda: Engine(plan)), ("table_gemm_fp32", lambda: Engine(plan, table_gemm=True))]
for v in a.variants.split(","):
    p_, s_ = v.split("x")
    def mk(p_=p_, s_=s_):
        os.environ["SDRX_TC_PRODUCTS"], os.environ["SDRX_TC_STEPS"] = p_, s_
        return Engine(plan, tensor_cores=True)
    engines.append((f"tensor_cores_bf16x{p_}_drain{int(s_) * 16}", mk))

lib = _lib.load()
L = plan.rate.block_len
nb = blocks[0].nbytes
ring = 56
d_ring = lib.sdrx_device_alloc(0, ring * nb)
for i in range(ring):
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d_ring + i * nb), blocks[i % len(blocks)].ctypes.data_as(C.c_void_p), nb))
res = {"workload": desc, "blocks": a.blocks, "bar": 1e-4, "open_channel_blocks": int(sum(int(w.sql.sum()) for w in want)),
       "squelch_decisions": int(sum(w.sql.size for w in want))}
for name, mk in engines:
    eng = mk()
    iq, au, mis, snr_d = [], [], 0, []
    for b, w in zip(blocks, want):
        g = eng.process(b)
        iq.append([rel_rms(g.ch_iq[c], w.ch_iq[c]) for c in range(a.channels)])
        mis += int(np.sum(g.sql != w.sql))
        ok = np.isfinite(w.snr) & np.isfinite(g.snr)
        snr_d.append(float(np.max(np.abs(g.snr[ok] - w.snr[ok]))))
        for c in range(a.channels):
            if np.any(w.audio[c] != 0.0) and np.any(g.audio[c] != 0.0):
                au.append(rel_rms(g.audio[c], w.audio[c]))
    iq = np.asarray(iq)
    per_ch = np.sqrt((iq ** 2).mean(axis=0))
    r = {"ch_iq_rel_rms_max": float(iq.max()), "ch_iq_rel_rms_median": float(np.median(iq)),
         "ch_iq_rel_rms_max_channels_ever_open": float(iq[:, ever_open].max()),
         "ch_iq_rel_rms_max_carrier_channels": float(iq[:, is_car].max()),
         "channel_blocks_above_1e-4": int((iq > 1e-4).sum()), "channel_blocks": int(iq.size),
         "audio_rel_rms_max": float(max(au)) if au else None, "squelch_decisions_differing": mis, "snr_abs_diff_max_db": max(snr_d)}
    for i in range(10):
        eng.push_device(d_ring + (i % ring) * nb, 2, L)
    eng.sync()
    n = 400
    t0 = time.perf_counter()
    for i in range(n):
        eng.push_device(d_ring + (i % ring) * nb, 2, L)
    eng.sync()
    dt = (time.perf_counter() - t0) / n
    eng.set_timing(True)
    ks = []
    for i in range(5):
        eng.push_device(d_ring + i * nb, 2, L); eng.sync()
        ks.append(eng.last_timing())
    eng.set_timing(False)
    r.update({"us_per_block": dt * 1e6, "channel_MS_per_s": a.channels * L / dt / 1e6,
              "channelizer_kernel_us": float(np.median([k[0] for k in ks])) * 1e3, "block_serialised_us": float(np.median([k[1] for k in ks])) * 1e3})
    res[name] = r
    print(name, json.dumps(r), flush=True)
    eng.close()
if a.out:
    json.dump(res, open(a.out, "w"), indent=1)

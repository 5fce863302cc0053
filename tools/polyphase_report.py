#!/usr/bin/env python
"""Opt-in polyphase / DFT channelizer (SDRX_FLAG_POLYPHASE) against the oracle and against the parity path:
per-channel error, squelch-decision agreement, throughput.   python tools/polyphase_report.py [--out file.json]"""
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
a = ap.parse_args()
plan, carriers, desc = wl.config3(a.channels)
blocks = wl.synth_blocks(carriers, a.blocks, 1)
orx = ob.OracleRx(product_tables(10000000), plan.offsets)
poly = Engine(plan, polyphase=True)
ref = Engine(plan)
tge = Engine(plan, table_gemm=True)
iq_p, iq_d, iq_t, sql_mis_p, sql_mis_t, sql_total, au_p, au_t, opened = [], [], [], 0, 0, 0, [], [], 0
snr_d = []
for b in blocks:
    w = orx.process(b, 2)
    g = poly.process(b)
    d = ref.process(b)
    tg = tge.process(b)
    iq_t.append([rel_rms(tg.ch_iq[c], w.ch_iq[c]) for c in range(a.channels)])
    sql_mis_t += int(np.sum(tg.sql != w.sql))
    for c in range(a.channels):
        if np.any(w.audio[c] != 0.0) and np.any(tg.audio[c] != 0.0):
            au_t.append(rel_rms(tg.audio[c], w.audio[c]))
    iq_p.append([rel_rms(g.ch_iq[c], w.ch_iq[c]) for c in range(a.channels)])
    iq_d.append([rel_rms(d.ch_iq[c], w.ch_iq[c]) for c in range(a.channels)])
    sql_mis_p += int(np.sum(g.sql != w.sql)); sql_total += w.sql.size; opened += int(w.sql.sum())
    ok = np.isfinite(w.snr) & np.isfinite(g.snr)
    snr_d.append(float(np.max(np.abs(g.snr[ok] - w.snr[ok]))))
    for c in range(a.channels):
        if np.any(w.audio[c] != 0.0) and np.any(g.audio[c] != 0.0):
            au_p.append(rel_rms(g.audio[c], w.audio[c]))
iq_p, iq_d, iq_t = np.asarray(iq_p), np.asarray(iq_d), np.asarray(iq_t)
car = sorted({c.offset for c in carriers})
is_car = np.asarray([o in car for o in plan.offsets])
res = {"workload": desc, "blocks": a.blocks,
       "polyphase_vs_oracle": {"ch_iq_rel_rms_max": float(iq_p.max()), "ch_iq_rel_rms_median": float(np.median(iq_p)),
                               "ch_iq_rel_rms_max_carrier_channels": float(iq_p[:, is_car].max()),
                               "ch_iq_rel_rms_max_noise_only_channels": float(iq_p[:, ~is_car].max()),
                               "audio_rel_rms_max": float(max(au_p)) if au_p else None,
                               "squelch_decisions_differing": sql_mis_p, "squelch_decisions": sql_total, "open_channel_blocks": opened,
                               "snr_abs_diff_max_db": max(snr_d)},
       "parity_path_vs_oracle": {"ch_iq_rel_rms_max": float(iq_d.max()), "ch_iq_rel_rms_median": float(np.median(iq_d))},
       "table_gemm_vs_oracle": {"ch_iq_rel_rms_max": float(iq_t.max()), "ch_iq_rel_rms_median": float(np.median(iq_t)),
                                "ch_iq_rel_rms_max_carrier_channels": float(iq_t[:, is_car].max()),
                                "channel_blocks_above_1e-4": int((iq_t > 1e-4).sum()), "channel_blocks": int(iq_t.size),
                                "audio_rel_rms_max": float(max(au_t)) if au_t else None,
                                "squelch_decisions_differing": sql_mis_t},
       "bar": 1e-4}
# ---- throughput: device-resident blocks, whole chain
lib = _lib.load()
L = plan.rate.block_len
nb = blocks[0].nbytes
ring = 56
d_ring = lib.sdrx_device_alloc(0, ring * nb)
for i in range(ring):
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d_ring + i * nb), blocks[i % len(blocks)].ctypes.data_as(C.c_void_p), nb))
for name, eng in (("polyphase", poly), ("parity_path", ref), ("table_gemm", tge)):
    for i in range(10):
        eng.push_device(d_ring + (i % ring) * nb, 2, L)
    eng.sync()
    t0 = time.perf_counter()
    n = 400
    for i in range(n):
        eng.push_device(d_ring + (i % ring) * nb, 2, L)
    eng.sync()
    dt = (time.perf_counter() - t0) / n
    eng.set_timing(True)
    eng.push_device(d_ring, 2, L); eng.sync()
    k_ms, blk_ms = eng.last_timing()
    eng.set_timing(False)
    res[name + "_throughput"] = {"us_per_block": dt * 1e6, "channel_MS_per_s": a.channels * L / dt / 1e6, "channelizer_kernel_us": k_ms * 1e3}
print(json.dumps(res, indent=1))
if a.out:
    json.dump(res, open(a.out, "w"), indent=1)

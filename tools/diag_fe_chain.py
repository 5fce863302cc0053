#!/usr/bin/env python
"""Diagnostic: where the time goes when the Airspy front end feeds the 960-channel chain (run on the GPU box)."""
import os, sys, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from sdrx_b200 import AirspyFrontEnd, ChannelPlan, Engine, PinnedBuffer, _lib
from tests.test_airspy_fe import adc_stream
from oracle import bindings as ob

N, STEPS = 640000, 40
offs = list(range(-480, 480))
eng = Engine(ChannelPlan.from_offsets(10000000, offs))
fe = AirspyFrontEnd(max_samples=N, packed=True, beside_engine=os.environ.get("FE_NARROW", "0") == "1")
blk = ob.airspy_pack12(adc_stream(N, seed=20))
pin = PinnedBuffer(blk.nbytes); v = pin.view(np.uint32, blk.size); v[:] = blk
st_fe = torch.cuda.ExternalStream(fe.stream)
st_en = torch.cuda.ExternalStream(eng.stream)
res = {}
d_static = _lib.load().sdrx_device_alloc(0, N // 2 * 8)
for mode in ("fe_only", "engine_only", "independent", "chain"):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(STEPS)]
    for i in range(6):          # warm-up of this mode (schedules, lazy loading)
        if mode == "fe_only": fe.process(v, fetch=False)
        elif mode == "engine_only": eng.push_device(fe.output_ptr, 2, N // 2)
        elif mode == "independent": fe.process(v, fetch=False); eng.push_device(d_static, 2, N // 2)
        else: fe.feed(eng, v)
    eng.sync(); fe.sync(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(STEPS):
        if mode == "fe_only":
            ev[i][0].record(st_fe); fe.process(v, fetch=False); ev[i][1].record(st_fe)
        elif mode == "engine_only":
            eng.push_device(fe.output_ptr, 2, N // 2)
        elif mode == "independent":
            ev[i][0].record(st_fe); fe.process(v, fetch=False); ev[i][1].record(st_fe); eng.push_device(d_static, 2, N // 2)
        else:
            ev[i][0].record(st_fe); fe.feed(eng, v); ev[i][1].record(st_fe)
    t_enq = (time.perf_counter() - t0) / STEPS * 1e6
    if mode != "fe_only":
        eng.set_timing(True)
        if mode == "engine_only": eng.push_device(fe.output_ptr, 2, N // 2)
        elif mode == "independent": fe.process(v, fetch=False); eng.push_device(d_static, 2, N // 2)
        else: fe.feed(eng, v)
        eng.sync(); fe.sync()
        bank_ms, block_ms = eng.last_timing()
        eng.set_timing(False)
    eng.sync(); fe.sync(); torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / STEPS * 1e6
    r = {"wall_us_per_step": dt, "host_enqueue_us_per_step": t_enq}
    if mode != "fe_only": r["last_bank_us"] = bank_ms * 1e3; r["last_block_us"] = block_ms * 1e3
    if mode != "engine_only":
        spans = [a.elapsed_time(b) * 1e3 for a, b in ev[5:]]
        gaps = [ev[i][0].elapsed_time(ev[i + 1][0]) * 1e3 for i in range(5, STEPS - 1)]
        r["fe_span_us_median"] = float(np.median(spans)); r["fe_period_us_median"] = float(np.median(gaps))
    res[mode] = r

# ---- one timeline on the GPU's own clock: the front end's five passes of call k+1 against the bank launch of block k
import ctypes as C
lib = _lib.load()
for _ in range(6):
    fe.feed(eng, v)
eng.sync(); fe.sync()
lib.sdrx_debug_bank_trace(1); lib.sdrx_debug_fe_trace(1)
for _ in range(6):
    fe.feed(eng, v)
# one more front-end call with no engine push behind it: it runs next to the LAST bank launch, whose trace is the one kept
ev_in = lib.sdrx_engine_ingest_event(eng.h)
lib.sdrx_airspy_fe_wait_event(fe.h, C.c_void_p(ev_in))
fe.process(v, fetch=False)
eng.sync(); fe.sync(); torch.cuda.synchronize()
bb = (C.c_uint64 * (160 * 16 * 2))(); nb = lib.sdrx_debug_bank_trace_read(bb, len(bb))
fb = (C.c_uint64 * 10)(); lib.sdrx_debug_fe_trace_read(fb)
lib.sdrx_debug_bank_trace(0); lib.sdrx_debug_fe_trace(0)
bt = np.frombuffer(bb, dtype=np.uint64)[:nb].reshape(-1, 16, 2).astype(np.int64)
bt = bt[bt[:, 0, 0] != 0]
b0, b1 = bt[:, :, 0].min(), bt[:, :, 1].max()
ft = np.frombuffer(fb, dtype=np.uint64).astype(np.int64).reshape(5, 2)
tl = {"bank": [0.0, (b1 - b0) / 1e3], "bank_ctas": int(len(bt)),
      "bank_cta_start_spread_us": float((bt[:, :, 0].min(axis=1).max() - b0) / 1e3)}
for name, (a, b) in zip(("fe_partial", "fe_scan", "fe_dc", "fe_iq", "fe_carry"), ft):
    tl[name] = [(a - b0) / 1e3, (b - b0) / 1e3]
res["timeline_us_relative_to_bank_start"] = tl
print(json.dumps(res, indent=1))

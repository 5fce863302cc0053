#!/usr/bin/env python
"""Ingest-stage measurement (SURVEY.md 8d "HBM roofline applies only to the ingest-only stage"): the Airspy real -> IQ
front end alone on device-resident buffers, and front end + whole chain fed from pinned host memory in the ADC's
packed 12-bit wire format.  Run on the GPU box:  python tools/bench_frontend.py > gpurun_out/frontend.json"""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                                   # noqa: E402  (events / streams only)
from sdrx_b200 import AirspyFrontEnd, ChannelPlan, Engine, PinnedBuffer, _lib   # noqa: E402
from tests.test_airspy_fe import adc_stream                    # noqa: E402
from oracle import bindings as ob                              # noqa: E402  (test-input packer only)

N = 640000                       # real samples per 32 ms block at 20 MS/s
STEPS = 50
out = {}
lib = _lib.load()
peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except OSError:
    pass
hbm = float(peaks.get("hbm_gbs", 6460.0))
for packed in (False, True):
    x = adc_stream(N, seed=3)
    raw = ob.airspy_pack12(x) if packed else x
    d = lib.sdrx_device_alloc(0, raw.nbytes)
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d), raw.ctypes.data_as(C.c_void_p), raw.nbytes))
    fe = AirspyFrontEnd(max_samples=N, packed=packed)
    st = torch.cuda.ExternalStream(fe.stream)
    for _ in range(5):
        fe.process_device(d, N)
    fe.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(STEPS):
        fe.process_device(d, N)
    e1.record(st)
    fe.sync()
    us = e0.elapsed_time(e1) * 1e3 / STEPS
    algo = raw.nbytes + N // 2 * 8                          # wire bytes in + float IQ out, each once
    out["packed12" if packed else "u16"] = {
        "us_per_block": us, "launches_per_block": 3, "algorithmic_bytes": algo, "achieved_gbs": algo / us / 1e3,
        "hbm_peak_gbs": hbm, "frac": algo / us / 1e3 / hbm,
        "real_msps": N / us, "note": "latency-bound: 5000 threads x 2176 sequential remove_dc steps (warm-up + segment)"}
    fe.close()
    lib.sdrx_device_free(0, C.c_void_p(d))

# chain: packed 12-bit ADC samples from pinned host memory -> front end -> 960-channel chain -> results to host
offs = list(range(-480, 480))
plan = ChannelPlan.from_offsets(10000000, offs)
eng = Engine(plan)
fe = AirspyFrontEnd(max_samples=N, packed=True, beside_engine=True)
blocks = [ob.airspy_pack12(adc_stream(N, seed=20 + i)) for i in range(4)]
pins = [PinnedBuffer(b.nbytes) for b in blocks]
for p, b in zip(pins, blocks):
    p.view(np.uint32, b.size)[:] = b
views = [p.view(np.uint32, b.size) for p, b in zip(pins, blocks)]
LAG = 3
osets = [eng.make_output_buffers(pinned=True) for _ in range(LAG + 1)]
for i in range(4):
    fe.feed(eng, views[i % 4])
    eng.outputs_into(osets[i % (LAG + 1)][0])
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(STEPS):
    fe.feed(eng, views[i % 4])
    eng.fetch_outputs(osets[i % (LAG + 1)][0])
    if i >= LAG:
        eng.wait_outputs(LAG)
for back in range(LAG - 1, -1, -1):
    eng.wait_outputs(back)
eng.sync()
dt = (time.perf_counter() - t0) / STEPS
out["chain_from_packed12"] = {"channels": len(offs), "ms_per_step": dt * 1e3, "channel_MSps": len(offs) * 320000 / dt / 1e6,
                              "h2d_bytes_per_step": int(blocks[0].nbytes), "d2h_bytes_per_step": osets[0][2],
                              "note": "wall clock, host buffers, 0.96 MB of wire per 32 ms block instead of 2.56 MB of float IQ"}
print(json.dumps(out, indent=1))

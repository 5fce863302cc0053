#!/usr/bin/env python
"""Diagnostic: per-CTA / per-warp timeline of one channel-bank launch (960 channels, 10 MS/s), from the kernel's own
globaltimer trace.  Run on the GPU box; prints where the launch's critical path is."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sdrx_b200 import ChannelPlan, Engine, _lib  # noqa: E402

lib = _lib.load()
offs = list(range(-480, 480))
eng = Engine(ChannelPlan.from_offsets(10000000, offs))
L = 320000
rng = np.random.Generator(np.random.PCG64(1))
blk = (rng.standard_normal(L) + 1j * rng.standard_normal(L)).astype(np.complex64) * 0.01
for _ in range(4):
    eng.process(blk, 2)
lib.sdrx_debug_bank_trace(1)
eng.process(blk, 2)
eng.sync()
buf = (C.c_uint64 * (160 * 16 * 2))()
n = lib.sdrx_debug_bank_trace_read(buf, len(buf))
lib.sdrx_debug_bank_trace(0)
t = np.frombuffer(buf, dtype=np.uint64)[:n].reshape(-1, 16, 2).astype(np.int64)
used = t[:, 0, 0] != 0
t = t[used]
t0 = t[:, :, 0].min()
start = (t[:, :, 0] - t0) / 1e3
end = (t[:, :, 1] - t0) / 1e3
cta_start, cta_end = start.min(axis=1), end.max(axis=1)
dur = cta_end - cta_start
order = np.argsort(cta_end)
res = {
    "ctas": int(used.sum()),
    "launch_span_us": float(cta_end.max()),
    "cta_start_us": {"min": float(cta_start.min()), "median": float(np.median(cta_start)), "max": float(cta_start.max())},
    "cta_end_us": {"min": float(cta_end.min()), "p10": float(np.percentile(cta_end, 10)), "median": float(np.median(cta_end)),
                   "p90": float(np.percentile(cta_end, 90)), "max": float(cta_end.max())},
    "cta_duration_us": {"min": float(dur.min()), "median": float(np.median(dur)), "max": float(dur.max())},
    "warp_end_spread_within_cta_us": {"median": float(np.median(end.max(axis=1) - end.min(axis=1))),
                                      "max": float((end.max(axis=1) - end.min(axis=1)).max())},
    "mean_busy_fraction_of_span": float(dur.sum() / (cta_end.max() * len(dur))),
    "slowest_ctas": [{"cta": int(i), "start": float(cta_start[i]), "end": float(cta_end[i])} for i in order[-8:][::-1]],
    "fastest_ctas": [{"cta": int(i), "start": float(cta_start[i]), "end": float(cta_end[i])} for i in order[:8]],
}
print(json.dumps(res, indent=1))
np.save(os.path.join(ROOT, "gpurun_out", "bank_trace.npy"), np.stack([start, end], axis=-1))
eng.close()

#!/usr/bin/env python
"""Per-kernel timeline of the engine's pipeline in steady state (sdrx_debug_timeline): which stream sets the step.
usage: python tools/diag_timeline.py [--open-all] [--blocks 12] [--out gpurun_out/timeline.json]"""
import argparse, ctypes as C, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sdrx_b200 import Engine, _lib, workloads as wl

ap = argparse.ArgumentParser()
ap.add_argument("--open-all", action="store_true")
ap.add_argument("--blocks", type=int, default=12)
ap.add_argument("--out", default=None)
ap.add_argument("--host", action="store_true", help="push from pinned host buffers (H2D inside the pipeline) instead of device-resident blocks")
ap.add_argument("--mode", default="bank", choices=["bank", "tgemm", "tc", "poly"])
a = ap.parse_args()
plan, carriers, desc = wl.config3(960, 0.0 if a.open_all else None)
blocks = wl.synth_blocks(carriers, 8, 1)
lib = _lib.load()
eng = Engine(plan, table_gemm=a.mode == "tgemm", tensor_cores=a.mode == "tc", polyphase=a.mode == "poly")
L = plan.rate.block_len
nb = blocks[0].nbytes
d = lib.sdrx_device_alloc(0, 8 * nb)
for i, b in enumerate(blocks):
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d + i * nb), b.ctypes.data_as(C.c_void_p), nb))
if a.host:
    from sdrx_b200 import PinnedBuffer
    pins = [PinnedBuffer(nb) for _ in blocks]
    hv = []
    for p_, b in zip(pins, blocks):
        v = p_.view(np.complex64, L); v[:] = b; hv.append(v)
    feed = lambda i: eng.push(hv[i % 8])
else:
    feed = lambda i: eng.push_device(d + (i % 8) * nb, 2, L)
for i in range(40):
    feed(i)
_lib.check(lib.sdrx_debug_timeline(eng.h, 1))
for i in range(40):
    feed(i)
out = np.zeros((16, 9), np.float32)
n = lib.sdrx_debug_timeline_read(eng.h, out.ctypes.data_as(_lib.c_float_p), a.blocks)
if n <= 0:
    print("timeline read failed:", lib.sdrx_last_error().decode(), file=sys.stderr)
names = ["start", "prep", "bank", "tail", "det", "agc", "env", "lfagc", "mix"]
t = out[:n] * 1e3
print(("open-all " if a.open_all else "") + "completion times, us after block start on the main stream; step = start-to-start")
print("blk   step " + " ".join(f"{x:>8s}" for x in names[1:]))
for i in range(n):
    step = t[i, 0] - t[i - 1, 0] if i else float("nan")
    print(f"{i:3d} {step:6.1f} " + " ".join(f"{t[i, k] - t[i, 0]:8.1f}" for k in range(1, 9)))
dur = {"bank": t[1:, 2] - np.maximum(t[1:, 0], t[1:, 1]), "tail": t[1:, 3] - t[1:, 2],
       "det": t[1:, 4] - np.maximum(t[1:, 3], t[:-1, 4]), "agc": t[1:, 5] - np.maximum(t[1:, 3], t[:-1, 7]),
       "env": t[1:, 6] - np.maximum(t[1:, 5], t[1:, 4]), "lfagc": t[1:, 7] - t[1:, 6], "mix": t[1:, 8] - np.maximum(t[1:, 7], t[:-1, 8])}
if a.mode == "tc":       # in this mode the engine marks the fold's completion in the `tail` slot (prep stream, behind k_prep)
    dur["tail"] = t[1:, 3] - t[1:, 1]
print("durations in the pipeline (completion minus the later of its stream predecessor and its dependency), us, median:")
print("  " + "  ".join(f"{k} {np.median(v):.1f}" for k, v in dur.items()) + f"  step {np.median(np.diff(t[:, 0])):.1f}")
if a.out:
    json.dump({"desc": desc, "open_all": a.open_all, "names": names, "completion_us": t.tolist(),
               "median_duration_us": {k: float(np.median(v)) for k, v in dur.items()}, "median_step_us": float(np.median(np.diff(t[:, 0])))},
              open(a.out, "w"))

#!/usr/bin/env python
"""Where the end-to-end step goes: the host loop of bench.py's e2e leg with parts left out.   python tools/e2e_probe.py [--mode tc]"""
import argparse, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sdrx_b200 import Engine, PinnedBuffer, workloads as wl

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="tc")
ap.add_argument("--steps", type=int, default=400)
ap.add_argument("--fmt", default="f32", choices=["f32", "s16"], help="wire format of the pushed blocks (s16: half the bytes across the link)")
a = ap.parse_args()
plan, carriers, _ = wl.config3(960)
blocks = wl.synth_blocks(carriers, 8, 1)
L = plan.rate.block_len
hv = []
if a.fmt == "s16":
    pins = [PinnedBuffer(b.nbytes // 2) for b in blocks]
    for p, b in zip(pins, blocks):
        v = p.view(np.int16, 2 * L); v[:] = np.clip(np.round(b.view(np.float32) * 32768.0), -32768, 32767).astype(np.int16); hv.append(v)
else:
    pins = [PinnedBuffer(b.nbytes) for b in blocks]
    for p, b in zip(pins, blocks):
        v = p.view(np.complex64, L); v[:] = b; hv.append(v)
eng = Engine(plan, tensor_cores=a.mode == "tc", polyphase=a.mode == "poly")
want = ("audio_rows", "audio_chan", "n_audio", "sql_state", "snr", "mix", "pcm", "pwr_dbfs")
LAG = 5
osets = [eng.make_output_buffers(pinned=True, want=want) for _ in range(LAG + 1)]

def run(fetch, wait, n=a.steps):
    for i in range(8):
        eng.push(hv[i % 8])
    eng.sync()
    t0 = time.perf_counter()
    for i in range(n):
        eng.push(hv[i % 8])
        if fetch:
            eng.fetch_outputs(osets[i % (LAG + 1)][0])
        if wait and i >= LAG:
            eng.wait_outputs(LAG)
    t1 = time.perf_counter()
    eng.sync()
    t2 = time.perf_counter()
    return (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6

for name, f, w in (("push only", False, False), ("push + fetch", True, False), ("push + fetch + wait(5)", True, True)):
    host, total = run(f, w)
    print(f"{name:26s} host issue {host:6.1f} us/step   with the final sync {total:6.1f} us/step", flush=True)

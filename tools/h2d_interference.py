#!/usr/bin/env python
"""Does host-to-device copy traffic slow the pipeline's kernels?  The device-resident pipeline (push_device) timed alone and with a
second CUDA stream copying 2.56 MB blocks from pinned host memory back to back beside it.   python tools/h2d_interference.py"""
import ctypes as C, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sdrx_b200 import Engine, _lib, workloads as wl

plan, carriers, _ = wl.config3(960)
blocks = wl.synth_blocks(carriers, 8, 1)
lib = _lib.load()
L, nb = plan.rate.block_len, blocks[0].nbytes
d = lib.sdrx_device_alloc(0, 8 * nb)
for i, b in enumerate(blocks):
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d + i * nb), b.ctypes.data_as(C.c_void_p), nb))
mode = sys.argv[1] if len(sys.argv) > 1 else "tc"
eng = Engine(plan, tensor_cores=mode == "tc", polyphase=mode == "poly")
hbuf = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
dbuf = torch.empty(nb, dtype=torch.uint8, device="cuda")
side = torch.cuda.Stream()

def run(n, copies):
    for i in range(20):
        eng.push_device(d + (i % 8) * nb, 2, L)
    eng.sync(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(n):
        if copies:
            with torch.cuda.stream(side):
                for _ in range(copies):
                    dbuf.copy_(hbuf, non_blocking=True)
        eng.push_device(d + (i % 8) * nb, 2, L)
    eng.sync()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    return (t1 - t0) / n * 1e6

for copies in (0, 1, 2):
    print(f"{mode}: {copies} x 2.56 MB H2D per block beside the pipeline: {run(600, copies):6.1f} us per block", flush=True)

"""Debug probe for multi-stream engines: python tools/ms_probe.py N_STREAMS [mode]   mode: host | device | s16"""
import ctypes as C, sys, numpy as np
sys.path.insert(0, '.')
from sdrx_b200 import Engine, _lib, workloads as wl
from sdrx_b200.source import quantize
NS = int(sys.argv[1]); mode = sys.argv[2] if len(sys.argv) > 2 else "host"
plan, car, _ = wl.config4_stream(0)
blk = wl.synth_blocks(car, 1, 1)[0]
lib = _lib.load()
eng = Engine(plan, n_streams=NS)
L = plan.rate.block_len
if mode == "s16":
    raw = np.stack([quantize(blk, 1)] * NS)
else:
    raw = np.stack([blk] * NS)
if mode == "device":
    d = lib.sdrx_device_alloc(0, raw.nbytes)
    _lib.check(lib.sdrx_device_upload(0, C.c_void_p(d), raw.ctypes.data_as(C.c_void_p), raw.nbytes))
for i in range(8):
    if mode == "device":
        eng.push_device(d, 2, L)
    else:
        eng.push(raw)
    print(NS, mode, "queued", i, flush=True)
eng.sync()
print("synced", flush=True)
o = eng.outputs(with_iq=False)
print("open", int(o.sql.sum()))
eng.close()

#!/usr/bin/env python
"""First light of the tensor-core table GEMM (SDRX_FLAG_TGEMM_TC): its channel IQ against the FP32 table GEMM and the
oracle on the 8-channel scenario, then on config 3 (960 channels).   python tools/tc_first_light.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sdrx_b200 import Engine, workloads as wl
from tests.harness import rel_rms

plan, carriers, desc = wl.config3(int(sys.argv[1]) if len(sys.argv) > 1 else 960)
blocks = wl.synth_blocks(carriers, 3, 1)
f32 = Engine(plan, table_gemm=True)
tc = Engine(plan, tensor_cores=True)
print("uses tensor cores:", tc.uses_tensor_cores, "pairs", tc.bank_pairs, flush=True)
for b in blocks:
    a = f32.process(b)
    g = tc.process(b)
    e = np.asarray([rel_rms(g.ch_iq[c], a.ch_iq[c]) for c in range(len(plan.offsets))])
    print(f"tc vs fp32 table GEMM: channel IQ rel rms max {e.max():.3e} median {np.median(e):.3e} worst channel {int(e.argmax())}; "
          f"sql equal {np.array_equal(g.sql, a.sql)}", flush=True)
    if e.max() > 1e-2:
        c = int(e.argmax())
        print("  got ", g.ch_iq[c][:4], "\n  want", a.ch_iq[c][:4])
        print("  channel 0 got", g.ch_iq[0][:3], "want", a.ch_iq[0][:3])

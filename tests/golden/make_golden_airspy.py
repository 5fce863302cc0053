#!/usr/bin/env python
"""Golden vectors of the Airspy real -> IQ front end, made HERE from the reference's own libairspy
iqconverter_float.c + filters.h (oracle/_ref/libairspy_ref.so, built by oracle/Makefile from the sources where
they lie under /root/reference).  The fixture travels to the GPU box, where /root/reference does not exist.

    python tests/golden/make_golden_airspy.py     ->  tests/golden/airspy_fe.npz
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import bindings as ob            # noqa: E402
from tests.test_airspy_fe import adc_stream  # noqa: E402

N, SEED = 262144, 2024
x = adc_stream(N, seed=SEED)
fe = ob.OracleAirspyFE(ob.ref_hb_kernel(), "ref")
y = np.concatenate([fe.process(x[i:i + 8192]) for i in range(0, N, 8192)])
out = os.path.join(ROOT, "tests", "golden", "airspy_fe.npz")
np.savez_compressed(out, n=np.int64(N), seed=np.int64(SEED), head=y[:4096], tail=y[-4096:],
                    sha256=np.str_(hashlib.sha256(y.tobytes()).hexdigest()))
print("wrote", out, os.path.getsize(out), "bytes")

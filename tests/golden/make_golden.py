#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE build (oracle/_ref/libsdrx_ref.so: the reference's
own msd.hpp / agc.hpp / fir.hpp / coeffs compiled with the pinned portable flags, see oracle/Makefile).

The reference ships no golden vectors for this path, so these outputs of the reference
itself, run in the build container, are the fixtures that pin the oracle and the CUDA path on the GPU
box, where /root/reference does not exist.  Run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from tests import scenarios as sc  # noqa: E402
from tests.harness import digest, run_cpu  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def emit(scn, keep_blocks):
    src = scn.source()
    blocks = [src.block(b) for b in range(scn.n_blocks)]
    outs = run_cpu(scn, "ref", blocks)
    d = {
        "input_sha256": np.array([digest(b) for b in blocks]),
        "sql": np.stack([o.sql for o in outs]),
        "snr": np.stack([o.snr for o in outs]),
        "pwr_dbfs": np.array([o.pwr_dbfs for o in outs], dtype=np.float32),
        "keep_blocks": np.array(keep_blocks, dtype=np.int32),
        "ch_iq": np.stack([outs[b].ch_iq for b in keep_blocks]),
        "audio": np.stack([outs[b].audio for b in keep_blocks]),
        "mix": np.stack([outs[b].mix for b in keep_blocks]),
        "pcm": np.stack([outs[b].pcm for b in keep_blocks]),
        "ch_iq_sha256": np.array([digest(o.ch_iq) for o in outs]),
        "pcm_sha256": np.array([digest(o.pcm) for o in outs]),
    }
    path = os.path.join(HERE, scn.name + ".npz")
    np.savez_compressed(path, **d)
    print(path, os.path.getsize(path), "bytes; squelch transitions:",
          int(np.sum(np.diff(d["sql"].astype(int), axis=0) != 0)))


if __name__ == "__main__":
    emit(sc.config1(), [0, 8, 9, 20, 28, 29])
    emit(sc.config2(), [0, 3, 4, 7, 11])
    emit(sc.config2(sc.FMT_S16, 4), [0, 3])
    emit(sc.centre_channel(), [0, 2])
    emit(sc.fm_scenarios()[0], [1, 3, 4, 6])      # FM + AM channels: open, noise while open, ramp-down, re-open

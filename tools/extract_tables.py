#!/usr/bin/env python
"""Extract the reference's coefficient tables as DATA into sdrx_b200/data/tables.npz.

SURVEY.md §2 #11: the decimation FIR taps (src/filters/*.hpp), the audio band-pass
`coeff_bp4am_channel`, `passband_shape`, `ramp_up`, `ramp_down` (src/coeffs.hpp) and the
per-rate N / z / stage lists (src/sdrx.cpp:1584-1677) are numeric tables the product consumes
verbatim.  They are read out of the reference's own headers through the compiled harness
(oracle/_ref/libsdrx_ref.so) -- nothing is retyped -- and stored as a binary archive that travels
to the GPU box, where /root/reference does not exist.

Run in the build container:  python tools/extract_tables.py
tests/test_tables.py re-checks the committed archive against the reference build when present.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import bindings as ob  # noqa: E402  (tooling, not product)


def main() -> None:
    out = {}
    for rate in ob.RATES:
        N, z, stages = ob.ref_rate_config(rate)
        out[f"rate_{rate}_N"] = np.int32(N)
        out[f"rate_{rate}_z"] = np.int32(z)
        out[f"rate_{rate}_m"] = np.asarray([s[0] for s in stages], dtype=np.int32)
        out[f"rate_{rate}_ntaps"] = np.asarray([len(s[1]) for s in stages], dtype=np.int32)
        out[f"rate_{rate}_taps"] = np.concatenate([s[1] for s in stages]).astype(np.float32)
    for name in ob.TABLE_NAMES:
        out[name] = ob.ref_table(name)
    # libairspy/libairspy/src/filters.h:28-30 HB_KERNEL_FLOAT: the half-band kernel of the real -> IQ front end
    out["airspy_hb_kernel"] = ob.ref_hb_kernel()
    path = os.path.join(ROOT, "sdrx_b200", "data", "tables.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Distance of the specialised channel bank from the CPU oracle on config 3 (960 channels), per channel, with
conjugate pairing on and off -- run on the GPU box; writes gpurun_out/parity_margins_pairing.json."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenarios as sc          # noqa: E402
from tests.harness import rel_rms, run_cpu, run_engine   # noqa: E402

scn = sc.config3(n_blocks=3)
src = scn.source()
blocks = [src.block(b) for b in range(scn.n_blocks)]
want = run_cpu(scn, "oracle", blocks)
iq_w = np.stack([o.ch_iq for o in want])
out = {"scenario": scn.name, "bar": 1e-4}
for name, kw in (("paired", {}), ("unpaired", {"no_pairing": True}), ("generic", {"force_generic": True})):
    got, _ = run_engine(scn, blocks=blocks, **kw)
    iq_g = np.stack([o.ch_iq for o in got])
    errs = np.asarray([rel_rms(iq_g[:, c], iq_w[:, c]) for c in range(iq_w.shape[1])])
    au = [rel_rms(np.stack([o.audio[c] for o in got]), np.stack([o.audio[c] for o in want]))
          for c in range(iq_w.shape[1]) if any(o.audio[c].any() for o in want)]
    out[name] = {"ch_iq_rel_rms_max": float(errs.max()), "ch_iq_rel_rms_median": float(np.median(errs)),
                 "audio_rel_rms_max": float(max(au)) if au else None,
                 "squelch_equal": bool(all(np.array_equal(g.sql, w.sql) for g, w in zip(got, want)))}
    print(name, out[name])
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "parity_margins_pairing.json"), "w") as f:
    json.dump(out, f, indent=1)

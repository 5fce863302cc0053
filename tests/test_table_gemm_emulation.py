"""The arithmetic of the table-GEMM channelizers against the oracle, emulated on the CPU (tools/emul_exact_gemm.py, numpy, no GPU):
the fold + GEMM over the reference's own translator tables is the reference's map -- evaluated exactly it differs from the oracle
only by the oracle's own FP32 rounding, far inside the 1e-4 bar -- and compensated addition of the 16-entry partial sums is never
worse than plain FP32 addition.  One block, the channels next to the strongest carriers plus a random sample."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))


def test_exact_table_gemm_is_the_reference_map(tmp_path):
    out = str(tmp_path / "emul.json")
    subprocess.run([sys.executable, os.path.join(HERE, "..", "tools", "emul_exact_gemm.py"), "--blocks", "1", "--out", out],
                   check=True, capture_output=True, timeout=600)
    r = json.load(open(out))
    assert r["exact"]["max"] < 5e-5 and r["exact"]["above_1e-4"] == 0, r["exact"]
    for k in ("fold32", "fold32_u16", "exact_u16", "exact_u16k"):
        assert r[k]["above_1e-4"] == 0, (k, r[k])
    assert r["exact_u16k"]["max"] <= r["exact_u16"]["max"] * 1.05, (r["exact_u16k"], r["exact_u16"])

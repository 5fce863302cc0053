"""The device log10f (sdrx_b200/csrc/common.cuh:log10f_glibc) restates glibc's algorithm so that the
squelch SNR is bit-identical to the reference's; its C twin is checked here against the host libm."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_log10f_restatement_matches_host_libm(tmp_path):
    exe = str(tmp_path / "log10f_check")
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(HERE, "aux", "log10f_check.c"), "-lm"],
                   check=True)
    out = subprocess.run([exe, "97"], check=True, capture_output=True, text=True).stdout
    assert "logf diff=0 log10f diff=0" in out, out


def test_device_source_carries_the_same_constants():
    src = open(os.path.join(HERE, "..", "sdrx_b200", "csrc", "common.cuh")).read()
    ref = open(os.path.join(HERE, "aux", "log10f_check.c")).read()
    import re
    consts = set(re.findall(r"-?0x1[.0-9a-f]*p[+-][0-9]+", ref))
    assert len(consts) >= 35
    for c in consts:
        assert c in src, c

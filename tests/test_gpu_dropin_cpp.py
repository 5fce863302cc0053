"""The C++ drop-in headers (include/sdrx_b200/{iqsample,msd,ds,agc,fir}.hpp) driven the way sdrx.cpp drives
the reference's classes (tests/cpp/dropin_main.cpp), compared with the oracle's restatement of those classes.
The CPU part only checks that the headers compile against include/sdrx_b200.h and link to the library."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "dropin_main.cpp")
BIN = os.path.join(ROOT, "tests", "cpp", "_build", "dropin_test")


def build_driver() -> str:
    os.makedirs(os.path.dirname(BIN), exist_ok=True)
    if os.path.exists(BIN) and os.path.getmtime(BIN) >= max(
            [os.path.getmtime(SRC), os.path.getmtime(os.path.join(ROOT, "include", "sdrx_b200.h"))] +
            [os.path.getmtime(os.path.join(ROOT, "include", "sdrx_b200", f))
             for f in os.listdir(os.path.join(ROOT, "include", "sdrx_b200"))]):
        return BIN
    from sdrx_b200 import build as b
    b.build()
    cmd = ["g++", "-std=c++20", "-O2", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), "-o", BIN, SRC,
           "-L" + os.path.join(ROOT, "sdrx_b200"), "-lsdrx_b200", "-Wl,-rpath," + os.path.join("$ORIGIN", "..", "..", "..", "sdrx_b200"),
           "-lpthread"]
    subprocess.run(cmd, check=True)
    return BIN


def test_dropin_headers_compile_and_link():
    assert os.path.exists(build_driver())


@pytest.mark.gpu
def test_dropin_classes_match_oracle(tmp_path):
    from oracle import bindings as ob
    from sdrx_b200 import config as cf
    from tests.harness import rel_rms
    exe = build_driver()
    rate = 10000000
    rc = cf.rate_config(rate)
    offs = [-360, -37, 0, 2, 250, 479]
    NB, L = 3, rc.block_len
    rng = np.random.default_rng(21)
    t = np.arange(NB * L)
    iq = 0.02 * (rng.standard_normal(NB * L) + 1j * rng.standard_normal(NB * L))
    iq += 0.2 * (1 + 0.5 * np.sin(2 * np.pi * 1000 * t / rate)) * np.exp(2j * np.pi * (((-37 * rc.z) * t) % rc.N) / rc.N)
    iq = iq.astype(np.complex64)
    fir2 = np.ascontiguousarray(cf.tables()["coeff_bp4am_channel"], np.float32)
    spec = tmp_path / "spec.bin"
    with open(spec, "wb") as f:
        np.asarray([len(offs), rc.N, len(rc.stages), L, NB], np.int32).tofile(f)
        np.asarray([s[0] for s in rc.stages], np.int32).tofile(f)
        np.asarray([len(s[1]) for s in rc.stages], np.int32).tofile(f)
        for s in rc.stages:
            np.asarray(s[1], np.float32).tofile(f)
        for o in offs:
            np.asarray([1 if o else 0], np.int32).tofile(f)
            (cf.make_translator(rc.N, rc.z, o) if o else np.zeros(rc.N, np.complex64)).tofile(f)
        np.asarray([fir2.size], np.int32).tofile(f)
        fir2.tofile(f)
        iq.tofile(f)
    out = tmp_path / "out.bin"
    r = subprocess.run([exe, str(spec), str(out)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    C_ = len(offs)
    raw = np.fromfile(out, np.uint8)
    pos = 0

    def take(dtype, n):
        nonlocal pos
        a = raw[pos:pos + n * np.dtype(dtype).itemsize].view(dtype)
        pos += a.nbytes
        return a
    msd = take(np.complex64, NB * C_ * 512).reshape(NB, C_, 512)
    ds = take(np.complex64, NB * C_ * 512).reshape(NB, C_, 512)
    agc = take(np.complex64, NB * 512).reshape(NB, 512)
    lf = take(np.float32, NB * 512).reshape(NB, 512)
    f2 = take(np.float32, NB * 1024).reshape(NB, 1024)
    f3 = take(np.complex64, NB * 512).reshape(NB, 512)
    batched = int(take(np.int32, 1)[0])
    assert pos == raw.size

    om = [ob.OracleMSD(cf.make_translator(rc.N, rc.z, o) if o else None, rc.stages) for o in offs]
    want = np.stack([np.stack([m.decimate(iq[b * L:(b + 1) * L]) for m in om]) for b in range(NB)])
    assert np.array_equal(msd, want), "MSD::decimate (one object per channel) is not bit-exact"
    assert batched == 1, "the DS bank did not take the batched engine"
    for c in range(C_):
        assert rel_rms(ds[:, c], want[:, c]) < 2e-6, (c, rel_rms(ds[:, c], want[:, c]))
    oa, ol = ob.OracleAGC(), ob.OracleLfAGC()
    oa.set(1.0, 0.01, 1.0, 300.0)
    ol.set(1.0, 0.01, 1.0, True)
    o2, o3 = ob.OracleFIR(2, fir2), ob.OracleFIR(3, fir2)
    o2.set_gain(3.0)
    for b in range(NB):
        g = oa.adjust(want[b, 0])
        assert np.array_equal(agc[b], g)
        env = (np.abs(g.real) + np.abs(g.imag)).astype(np.float32)
        l_ = ol.adjust(env)
        assert np.array_equal(lf[b], l_)
        st = np.empty(1024, np.float32)
        st[0::2], st[1::2] = np.float32(0.8) * l_, np.float32(0.2) * l_
        assert np.array_equal(f2[b], o2.filter(st))
        assert np.array_equal(f3[b], o3.filter(want[b, 0]))

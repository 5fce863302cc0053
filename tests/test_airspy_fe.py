"""Airspy real -> IQ front end.

CPU: the plain-C restatement (oracle/airspy_fe_oracle.c) is pinned bit-exact against the reference's own
libairspy iqconverter_float.c + filters.h compiled here (oracle/_ref/libairspy_ref.so), and against the committed
golden vectors made from that build (tests/golden/make_golden_airspy.py).
GPU (-m gpu): the CUDA front end through the C ABI against the oracle -- bit-exact with SDRX_FE_EXACT_DC, and for the
default segment-parallel remove_dc bit-exact on these inputs too (reported), held to 1e-6 absolute; then the whole
chain fed from the front end against the oracle chain fed by the oracle front end."""
import os
import re

import numpy as np
import pytest

from oracle import bindings as ob
from sdrx_b200 import config as cf

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "airspy_fe.npz")
needs_ref = pytest.mark.skipif(not (ob.airspy_ref_available() or os.path.isdir(ob.REF_SRC)),
                               reason="oracle/_ref/libairspy_ref.so not available")


def adc_stream(n: int, seed: int = 7, dc: float = 37.0) -> np.ndarray:
    """12-bit real ADC samples at 20 MS/s: a few carriers around fs/4 (= the IQ band centre), DC offset, noise."""
    rng = np.random.Generator(np.random.PCG64(seed))
    t = np.arange(n, dtype=np.float64)
    x = np.zeros(n)
    for f, a in ((0.25 + 0.013, 500.0), (0.25 - 0.08, 200.0), (0.25 + 0.19, 90.0), (0.031, 40.0)):
        x += a * np.cos(2 * np.pi * f * t + rng.uniform(0, 6.28))
    x += rng.normal(0.0, 12.0, n) + 2048.0 + dc
    return np.clip(np.rint(x), 0, 4095).astype(np.uint16)


def hb():
    return np.ascontiguousarray(cf.tables()["airspy_hb_kernel"], np.float32)


@needs_ref
def test_extracted_half_band_kernel_matches_reference_header():
    assert np.array_equal(hb(), ob.ref_hb_kernel())
    assert hb().size == 47 and hb()[23] == np.float32(0.5)


@needs_ref
@pytest.mark.parametrize("chunks", [[65536], [4, 8, 400, 4000, 12, 61112], [32768, 32768], [8] * 40])
def test_restatement_is_bit_exact_with_libairspy(chunks):
    x = adc_stream(sum(chunks))
    a, b = ob.OracleAirspyFE(hb()), ob.OracleAirspyFE(hb(), "ref")
    pos = 0
    for n in chunks:
        assert np.array_equal(a.process(x[pos:pos + n]), b.process(x[pos:pos + n])), (pos, n)
        pos += n
    a.reset(), b.reset()
    assert np.array_equal(a.process(x[:4096]), b.process(x[:4096]))


def test_unpack12_round_trip_and_layout():
    x = adc_stream(4096, seed=3)
    fe = ob.OracleAirspyFE(hb())
    assert np.array_equal(fe.unpack12(ob.airspy_pack12(x)), x)
    # airspy.c:329-336: sample 0 is the top 12 bits of word 0, sample 7 the low 12 bits of word 2
    w = ob.airspy_pack12(np.asarray([0xABC, 0, 0, 0, 0, 0, 0, 0x123], np.uint16))
    assert w[0] >> 20 == 0xABC and w[2] & 0xfff == 0x123


def test_oracle_reproduces_golden_vectors():
    z = np.load(GOLDEN)
    x = adc_stream(int(z["n"]), seed=int(z["seed"]))
    fe = ob.OracleAirspyFE(hb())
    got = np.concatenate([fe.process(x[i:i + 8192]) for i in range(0, x.size, 8192)])
    assert np.array_equal(got[:z["head"].size], z["head"]) and np.array_equal(got[-z["tail"].size:], z["tail"])
    import hashlib
    assert hashlib.sha256(got.tobytes()).hexdigest() == str(z["sha256"])


def test_front_end_image_rejection_and_dc():
    """Property check of the restated algorithm: a tone at fs/4 + d lands at -2d (of the IQ rate) in the output, DC is removed."""
    n = 1 << 16
    t = np.arange(n)
    d = 0.02
    x = np.clip(np.rint(2048 + 100 + 900 * np.cos(2 * np.pi * (0.25 + d) * t)), 0, 4095).astype(np.uint16)
    y = ob.OracleAirspyFE(hb()).process(x)[4096:]
    spec = np.abs(np.fft.fft(y * np.hanning(y.size)))
    k = int(np.argmax(spec))
    f = k / y.size if k < y.size // 2 else k / y.size - 1.0
    assert abs(f + 2 * d) < 1e-3                           # IQ rate is half the real rate; the rotation mirrors the band
    assert abs(np.mean(y)) < 1e-3                          # the 100-count DC offset is gone


def test_library_exports_every_symbol_the_header_declares():
    from sdrx_b200 import _lib
    lib = _lib.load()
    text = open(os.path.join(ROOT, "include", "sdrx_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = set(re.findall(r"\b(sdrx_[a-z0-9_]+)\s*\(", text))
    assert len(names) > 50
    declared = {p[0] for p in _lib.PROTOTYPES}
    for n in sorted(names):
        assert hasattr(lib, n), f"{n} is declared in include/sdrx_b200.h but not exported"
        assert n in declared, f"{n} has no ctypes prototype in sdrx_b200/_lib.py"


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("packed", [False, True], ids=["u16", "packed12"])
@pytest.mark.parametrize("exact", [True, False, "narrow"], ids=["exact_dc", "parallel_dc", "narrow_dc"])
def test_cuda_front_end_against_oracle(packed, exact):
    from sdrx_b200 import AirspyFrontEnd
    narrow = exact == "narrow"                  # SDRX_FE_BESIDE_ENGINE: 512-sample segments staged through shared memory
    exact = exact is True
    chunks = [8, 640, 65536, 4096, 200000, 24, 640000, 1032, 520, 512, 1536] if not exact else [8, 640, 65536, 4096, 24]
    x = adc_stream(sum(chunks), seed=11)
    o = ob.OracleAirspyFE(hb())
    fe = AirspyFrontEnd(max_samples=640000, packed=packed, exact_dc=exact, beside_engine=narrow)
    pos, worst, differing, total = 0, 0.0, 0, 0
    for n in chunks:
        raw = x[pos:pos + n]
        want = o.process(raw)
        got = fe.process(ob.airspy_pack12(raw) if packed else raw)
        pos += n
        if exact:
            assert np.array_equal(got, want), (pos, n)
        worst = max(worst, float(np.max(np.abs(got - want))))
        differing += int(np.count_nonzero(got != want))
        total += want.size
    print(f"front end ({'packed' if packed else 'u16'}, {'exact' if exact else 'narrow' if narrow else 'parallel'} dc): "
          f"{differing} of {total} IQ samples differ from the oracle, max abs {worst:.3e}")
    assert worst < 1e-6                                    # full scale is 1.0
    fe.reset()
    o.reset()
    assert np.max(np.abs(fe.process(ob.airspy_pack12(x[:4096]) if packed else x[:4096]) - o.process(x[:4096]))) < 1e-6
    fe.close()


@pytest.mark.gpu
@pytest.mark.parametrize("beside", [False, True], ids=["wide", "beside_engine"])
def test_front_end_feeds_the_channel_bank(beside):
    """20 MS/s real ADC samples -> front end -> engine (device to device) == oracle front end -> oracle chain."""
    from sdrx_b200 import AirspyFrontEnd, ChannelPlan, Engine
    from tests.harness import AUDIO_TOL, product_tables, rel_rms
    rate, L = 10000000, 320000
    offs = [-300, -37, 2, 121, 400]
    n_blocks = 4
    rng = np.random.Generator(np.random.PCG64(5))
    t = np.arange(2 * L * n_blocks, dtype=np.float64)
    x = rng.normal(0.0, 6.0, t.size) + 2048 + 21
    for k, off in enumerate(offs):          # AM carriers at fs/4 - offset * 8.333 kHz: libairspy's fs/4 rotation mirrors the band
        f = 0.25 - off * (25000.0 / 3) / 20e6
        x += (60.0 + 25 * k) * (1 + 0.5 * np.cos(2 * np.pi * (700.0 + 300 * k) / 20e6 * t)) * np.cos(2 * np.pi * f * t)
    x = np.clip(np.rint(x), 0, 4095).astype(np.uint16)
    ofe = ob.OracleAirspyFE(hb())
    orx = ob.OracleRx(product_tables(rate), offs)
    fe = AirspyFrontEnd(max_samples=2 * L, beside_engine=beside)
    eng = Engine(ChannelPlan.from_offsets(rate, offs))
    opened = False
    for b in range(n_blocks):
        raw = x[2 * L * b: 2 * L * (b + 1)]
        want = orx.process(ofe.process(raw), 2)
        fe.feed(eng, raw)
        got = eng.outputs()
        assert np.array_equal(got.sql, want.sql)
        opened |= bool(want.sql.any())
        assert rel_rms(eng.channel_iq(), want.ch_iq) < AUDIO_TOL
        if np.any(want.mix):
            assert rel_rms(got.mix, want.mix) < AUDIO_TOL
    assert opened
    fe.close()
    eng.close()

"""GPU tests of the single-object class surface (msd.hpp / agc.hpp / fir.hpp mirrors) against the
oracle's restatement of the same classes -- these read like unit tests of the reference's classes."""
import numpy as np
import pytest

from oracle import bindings as ob
from sdrx_b200 import config as cf
from sdrx_b200.objects import AGC, FIR, FIR2, FIR3, MSD, LfAGC

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rate,off,ftfir", [(2400000, -3, False), (10000000, 7, False), (10000000, 0, False),
                                            (2560000, 11, False), (10000000, -233, True), (2400000, 5, True),
                                            (10000000, 0, True)])
def test_msd_any_chunking_and_copy(rate, off, ftfir):
    rc = cf.rate_config(rate)
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(50000) + 1j * rng.standard_normal(50000)).astype(np.complex64)
    tr = cf.make_translator(rc.N, rc.z, off) if off else None
    a, o = MSD(tr, rc.stages, use_ftfir=ftfir), ob.OracleMSD(tr, rc.stages, use_ftfir=ftfir)   # msd.hpp:54
    assert a.m() == o.m == rc.M
    pos = 0
    for n in [1, 3, 17, 1000, 4096, 7, 20001, 24875]:
        assert np.array_equal(a.decimate(x[pos:pos + n]), o.decimate(x[pos:pos + n])), n
        pos += n
    a2 = a.clone()
    more = (rng.standard_normal(5000) + 1j * rng.standard_normal(5000)).astype(np.complex64)
    want = o.decimate(more)
    assert np.array_equal(a.decimate(more), want) and np.array_equal(a2.decimate(more), want)
    # growing past the first capacity keeps the delay lines
    big = (rng.standard_normal(200000) + 1j * rng.standard_normal(200000)).astype(np.complex64)
    assert np.array_equal(a.decimate(big), o.decimate(big))
    a.close()
    a2.close()


def test_agc_and_lfagc_match_oracle_bit_for_bit():
    rng = np.random.default_rng(9)
    x = ((rng.standard_normal(4000) + 1j * rng.standard_normal(4000)) * np.repeat([0.01, 2.0, 0.0005, 0.3], 1000)).astype(np.complex64)
    a, o = AGC(), ob.OracleAGC()
    for lo, hi in ((0, 1500), (1500, 1501), (1501, 4000)):
        assert np.array_equal(a.adjust(x[lo:hi]), o.adjust(x[lo:hi]))
    assert a.gain() == o.gain
    a, o = AGC(), ob.OracleAGC()
    a.setReference(1.0), a.setAttack(1.0), a.setDecay(0.01), a.setMaxGain(300.0)     # sdrx.cpp:1702-1705
    o.set(1.0, 0.01, 1.0, 300.0)
    assert np.array_equal(a.adjust(x), o.adjust(x)) and a.gain() == o.gain
    s = np.abs(x).astype(np.float32)
    for active in (False, True):
        a, o = LfAGC(), ob.OracleLfAGC()
        a.setReference(1.0), a.setAttack(1.0), a.setDecay(0.01)
        (a.activate if active else a.deactivate)()
        o.set(1.0, 0.01, 1.0, active)
        assert np.array_equal(a.adjust(s), o.adjust(s)) and a.gain() == o.gain


def test_fir_classes_match_oracle_bit_for_bit():
    rng = np.random.default_rng(11)
    taps = cf.tables()["coeff_bp4am_channel"]
    s = rng.standard_normal(4000).astype(np.float32)
    x = (rng.standard_normal(4000) + 1j * rng.standard_normal(4000)).astype(np.complex64)
    for cls, kind, data in ((FIR, 1, s), (FIR2, 2, s), (FIR3, 3, x)):
        a, o = cls(taps), ob.OracleFIR(kind, taps)
        for lo, hi in ((0, 1000), (1000, 1002), (1002, 4000)):
            assert np.array_equal(a.filter(data[lo:hi]), o.filter(data[lo:hi])), (kind, lo)
        a.setGain(3.5)
        o.set_gain(3.5)
        assert np.array_equal(a.filter(data[:512]), o.filter(data[:512]))
        # a copy carries the delay line and the gain (the reference's implicit copy, fir.hpp:69-74, :176-182): both
        # continue the stream exactly like the oracle, which kept running
        b = a.copy()
        want = o.filter(data[512:1024])
        assert np.array_equal(b.filter(data[512:1024]), want), (kind, "copy")
        assert np.array_equal(a.filter(data[512:1024]), want), (kind, "original after the copy")
        assert b.gain() == 3.5
        b.close()
        a.close()

"""CPU tests of the oracle (test infrastructure): the plain-C restatement is pinned bit-exact against
the reference's own headers compiled here (oracle/_ref, skipped where /root/reference is absent and no
prebuilt library travelled) and against the committed golden vectors generated from that build."""
import os

import numpy as np
import pytest

from oracle import bindings as ob
from sdrx_b200 import config as cf
from tests import scenarios as sc
from tests.harness import digest, product_tables, run_cpu

HAVE_REF = ob.backend_available("ref") or os.path.isdir(ob.REF_SRC)
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="reference build oracle/_ref/libsdrx_ref.so not available")
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _assert_blocks_equal(a, b):
    for i, (x, y) in enumerate(zip(a, b)):
        assert x.pwr_dbfs == y.pwr_dbfs, i
        assert np.array_equal(x.ch_iq, y.ch_iq), f"block {i} channel IQ"
        assert np.array_equal(x.audio, y.audio), f"block {i} audio"
        assert np.array_equal(x.sql, y.sql), f"block {i} squelch"
        assert np.array_equal(x.snr, y.snr, equal_nan=True), f"block {i} snr"
        assert np.array_equal(x.mix_pre, y.mix_pre) and np.array_equal(x.mix, y.mix), f"block {i} mix"
        assert np.array_equal(x.pcm, y.pcm), f"block {i} pcm"


def test_fm_discriminator_recovers_the_tone():
    """Demod::demod FM (sdrx.cpp:181-194): a carrier with peak deviation d and tone f demodulates to a sine of
    amplitude 2 pi d / 16000 at f; the block after the carrier drops (squelch still open) is full-scale phase noise;
    the ramp-down block after that is the envelope, not the angle (sdrx.cpp:500)."""
    scn = sc.fm_scenarios()[0]
    outs = run_cpu(scn, "oracle")
    au = np.stack([o.audio for o in outs])
    sql = np.stack([o.sql for o in outs])
    for c, (dev, tone, blk) in {0: (2500.0, 800.0, 2), 2: (1200.0, 1500.0, 4)}.items():
        x = au[blk, c].astype(np.float64)
        assert sql[blk - 1, c] == 1 and sql[blk, c] == 1
        spec = np.abs(np.fft.rfft(x * np.hanning(512)))
        assert abs(np.argmax(spec) * (16000.0 / 512) - tone) <= 16000.0 / 512
        assert abs(np.max(np.abs(x)) - 2 * np.pi * dev / 16000.0) < 0.03
    assert sql[2, 0] == 1 and sql[3, 0] == 0               # carrier 0 is keyed off from block 3
    assert np.max(np.abs(au[3, 0])) > 3.0                  # demodulated while still open: noise over the full +-pi
    assert 0.0 < np.max(np.abs(au[4, 0])) < 0.05           # ramp-down: the (small) envelope
    assert np.all(au[5, 0] == 0.0)


@needs_ref
@pytest.mark.parametrize("scn", [sc.config1(), sc.config2(n_blocks=6), sc.config2(sc.FMT_S16, 3), sc.centre_channel()]
                         + sc.config2_variants() + sc.other_rates() + sc.ftfir_scenarios() + sc.fm_scenarios(),
                         ids=lambda s: s.name)
def test_restatement_is_bit_exact_with_reference_build(scn):
    src = scn.source()
    blocks = [src.block(b) for b in range(scn.n_blocks)]
    _assert_blocks_equal(run_cpu(scn, "oracle", blocks), run_cpu(scn, "ref", blocks))


@pytest.mark.parametrize("scn,name", [(sc.config1(), None), (sc.config2(), None), (sc.config2(sc.FMT_S16, 4), None),
                                      (sc.centre_channel(), None), (sc.fm_scenarios()[0], None)],
                         ids=lambda v: getattr(v, "name", ""))
def test_oracle_reproduces_golden_vectors(scn, name):
    z = np.load(os.path.join(GOLDEN, scn.name + ".npz"))
    src = scn.source()
    blocks = [src.block(b) for b in range(scn.n_blocks)]
    assert [digest(b) for b in blocks] == list(z["input_sha256"]), "synthetic source is not reproducing the golden input"
    outs = run_cpu(scn, "oracle", blocks)
    assert np.array_equal(np.stack([o.sql for o in outs]), z["sql"])
    assert np.array_equal(np.stack([o.snr for o in outs]), z["snr"], equal_nan=True)
    assert [digest(o.ch_iq) for o in outs] == list(z["ch_iq_sha256"])
    assert [digest(o.pcm) for o in outs] == list(z["pcm_sha256"])
    for j, b in enumerate(z["keep_blocks"]):
        assert np.array_equal(outs[b].ch_iq, z["ch_iq"][j])
        assert np.array_equal(outs[b].audio, z["audio"][j])
        assert np.array_equal(outs[b].mix, z["mix"][j])
        assert np.array_equal(outs[b].pcm, z["pcm"][j])


@needs_ref
def test_extracted_tables_match_reference_headers():
    for rate in ob.RATES:
        N, z, stages = ob.ref_rate_config(rate)
        rc = cf.rate_config(rate)
        assert (rc.N, rc.z, len(rc.stages)) == (N, z, len(stages))
        for (m0, h0), (m1, h1) in zip(rc.stages, stages):
            assert m0 == m1 and np.array_equal(h0, h1)
    t = cf.tables()
    for name in ob.TABLE_NAMES:
        assert np.array_equal(t[name], ob.ref_table(name)), name


def test_stage_tables_have_the_documented_shape():
    # 10 MS/s (m,H) = (5,17),(5,19),(5,21),(5,67); 2.4 MS/s (2,9),(3,13),(5,19),(5,45)
    assert [(m, len(h)) for m, h in cf.rate_config(10000000).stages] == [(5, 17), (5, 19), (5, 21), (5, 67)]
    assert [(m, len(h)) for m, h in cf.rate_config(2400000).stages] == [(2, 9), (3, 13), (5, 19), (5, 45)]
    for rate in cf.SAMPLE_RATES:
        rc = cf.rate_config(rate)
        assert rc.M * 16000 == rate
        for m, h in rc.stages:
            assert len(h) % 2 == 1 and np.allclose(h, h[::-1]) and abs(float(np.sum(h.astype(np.float64))) - 1.0) < 1e-3
    t = cf.tables()
    assert t["coeff_bp4am_channel"].size == 197 and t["passband_shape"].size == 512
    assert t["ramp_up"].size == 512 and t["ramp_down"].size == 512


@needs_ref
def test_translator_window_and_pan_positions_match_reference():
    for N, z, off in [(1200, 1, 479), (1200, 1, -480), (288, 1, -3), (576, 5, 17), (1728, 10, -33), (1536, 5, 60)]:
        assert np.array_equal(ob.translator(N, z, off, "oracle"), ob.translator(N, z, off, "ref"))
    assert np.array_equal(ob.window("oracle"), ob.window("ref"))
    for n in range(1, 33):
        assert [ob.audio_pos(c, n, "oracle") for c in range(n)] == [ob.audio_pos(c, n, "ref") for c in range(n)]


def test_pan_positions_of_eight_channels():
    # SURVEY.md §8d config 2
    assert [ob.audio_pos(c, 8) for c in range(8)] == [-2, -2, -1, -1, 1, 1, 2, 2]


@needs_ref
@pytest.mark.parametrize("rate", [2400000, 10000000, 2560000])
def test_msd_any_chunking_and_copy_semantics(rate):
    """MSD::decimate takes any in_len with state carried (msd.hpp:75); MSD is copyable (ds.hpp:34)."""
    N, z, stages = ob.ref_rate_config(rate)
    rng = np.random.default_rng(5)
    x = (rng.standard_normal(50000) + 1j * rng.standard_normal(50000)).astype(np.complex64)
    tr = ob.translator(N, z, 7, "ref")
    a, b = ob.OracleMSD(tr, stages, backend="oracle"), ob.OracleMSD(tr, stages, backend="ref")
    assert a.m == b.m
    pos, outs_a, outs_b = 0, [], []
    for n in [1, 3, 17, 1000, 4096, 7, 20001, 24875]:
        outs_a.append(a.decimate(x[pos:pos + n]))
        outs_b.append(b.decimate(x[pos:pos + n]))
        pos += n
    assert pos == x.size
    ya, yb = np.concatenate(outs_a), np.concatenate(outs_b)
    assert np.array_equal(ya, yb)
    whole = ob.OracleMSD(tr, stages, backend="oracle").decimate(x)
    assert np.array_equal(whole, ya)
    # a copy continues independently from the same state
    a2 = a.clone()
    more = (rng.standard_normal(5000) + 1j * rng.standard_normal(5000)).astype(np.complex64)
    assert np.array_equal(a.decimate(more), a2.decimate(more))


@needs_ref
def test_agc_lfagc_fir_classes_match_reference():
    rng = np.random.default_rng(9)
    x = ((rng.standard_normal(4000) + 1j * rng.standard_normal(4000)) * np.repeat([0.01, 2.0, 0.0005, 0.3], 1000)).astype(np.complex64)
    for be_args in [dict(), dict(attack=1.0, decay=0.01, reference=1.0, max_gain=300.0)]:
        a, b = ob.OracleAGC(backend="oracle"), ob.OracleAGC(backend="ref")
        if be_args:
            a.set(**be_args)
            b.set(**be_args)
        assert np.array_equal(a.adjust(x), b.adjust(x)) and a.gain == b.gain
    s = np.abs(x).astype(np.float32)
    for active in (False, True):
        a, b = ob.OracleLfAGC(backend="oracle"), ob.OracleLfAGC(backend="ref")
        a.set(1.0, 0.01, 1.0, active)
        b.set(1.0, 0.01, 1.0, active)
        assert np.array_equal(a.adjust(s), b.adjust(s)) and a.gain == b.gain
    taps = cf.tables()["coeff_bp4am_channel"]
    for kind, data in ((1, s), (2, s), (3, x)):
        a, b = ob.OracleFIR(kind, taps, "oracle"), ob.OracleFIR(kind, taps, "ref")
        a.set_gain(3.5)
        b.set_gain(3.5)
        for lo, hi in ((0, 1000), (1000, 1002), (1002, 4000)):
            assert np.array_equal(a.filter(data[lo:hi]), b.filter(data[lo:hi]))


def test_fft512_against_numpy_double():
    """The squelch FFT boundary is 'parity unpinned' (FFTW3f is not vendored): the oracle's FP32
    radix-2 FFT is checked against numpy's double-precision FFT instead."""
    rng = np.random.default_rng(3)
    for scale in (1.0, 1e-3):
        x = (scale * (rng.standard_normal(512) + 1j * rng.standard_normal(512))).astype(np.complex64)
        ref = np.fft.fft(x.astype(np.complex128))
        got = ob.fft512(x)
        assert np.linalg.norm(got - ref) / np.linalg.norm(ref) < 5e-7
    imp = np.zeros(512, np.complex64)
    imp[1] = 1.0
    assert np.allclose(ob.fft512(imp), np.exp(-2j * np.pi * np.arange(512) / 512), atol=1e-6)


def test_channel_name_helpers_agree_between_oracle_and_host_mirror():
    names = ["118.000", "118.005", "118.010", "118.015", "118.025", "118.030", "118.035", "118.040", "118.050",
             "118.055", "118.060", "118.065", "118.075", "118.080", "118.085", "118.090", "132.990", "121.500",
             "118.275", "127.415"]
    for n in names:
        assert cf.parse_fq(n, True) == ob.parse_fq(n, True) != 0
        for tuner in (118300000, 122000000, 127400000, 136900000):
            assert cf.channel_to_offset(n, tuner) == ob.channel_to_offset(n, tuner)
    for bad in ["118.020", "118.5", "abc.def", "118", "1.000", "118.0000"]:
        assert cf.parse_fq(bad, True) == ob.parse_fq(bad, True) == 0
    assert cf.parse_fq("433.92", False) == ob.parse_fq("433.92", False) == 433920000
    # 118.275 -> tuner 118.3 MHz, offset -3
    assert cf.tuner_fq_for(["118.275"]) == 118300000 and cf.channel_to_offset("118.275", 118300000) == -3
    for n in range(1, 40):
        assert [cf.get_audio_pos(c, n) for c in range(n)] == [ob.audio_pos(c, n) for c in range(n)]
    plan = cf.ChannelPlan.from_names(10000000, ["118.105/12", "121.500", "121.500", "125.230/0/AM"])
    assert plan.names == ["118.105", "121.500", "125.230"] and plan.sql_levels == [12.0, 9.0, 0.0]
    assert plan.tuner_fq == 121700000
    assert plan.modulation == ["AM", "AM", "AM"] and plan.modulation_codes() == [0, 0, 0]
    # NNN.NNN/SQL/AM|FM and --modulation as the default for channels without a suffix (sdrx.cpp:1197, :1351-1366)
    fm = cf.ChannelPlan.from_names(10000000, ["118.105/12/FM", "121.500", "125.230/0/AM"], default_mod="FM")
    assert fm.modulation == ["FM", "FM", "AM"] and fm.modulation_codes() == [1, 1, 0]
    assert cf.ChannelPlan.from_offsets(10000000, [1, 2]).modulation_codes() == [0, 0]
    with pytest.raises(ValueError):
        cf.ChannelPlan.from_offsets(10000000, [1, 2], modulation=["FM"]).modulation_codes()
    for off in range(-480, 480, 11):
        assert cf.channel_to_offset(cf.offset_to_channel(off, 122000000), 122000000) == off

"""Capture replay: rtl_sdr raw u8 IQ, airspy_rx raw / WAV INT16_IQ and FLOAT32_IQ, and
airspy_rx's real sample types for the front end.

CPU: the files FileSource reads back are the blocks that were written, header fields follow airspy_rx.c:126-190, the
int16 / u8 conventions are the reference's (rtl_dev.cpp:320-323, airspy.c:58-62).
GPU (-m gpu): a capture replayed through the engine gives what the oracle gives on the same file."""
import os
import struct

import numpy as np
import pytest

from oracle import bindings as ob
from sdrx_b200 import FMT_F32, FMT_S16, FMT_U8, Carrier, FileSource, SyntheticSource
from sdrx_b200 import source as so

RATE = 2400000


def band(fmt, n_blocks=3, rate=RATE):
    return SyntheticSource(rate, [Carrier(-3, 0.4, 1000.0, 0.5), Carrier(7, 0.1, 600.0, 0.3)], 0.01, 99, fmt, n_blocks)


@pytest.mark.parametrize("fmt", [FMT_U8, FMT_S16, FMT_F32], ids=["u8", "s16", "f32"])
@pytest.mark.parametrize("wav", [False, True], ids=["raw", "wav"])
def test_capture_round_trip(tmp_path, fmt, wav):
    src = band(fmt)
    p = str(tmp_path / "cap.bin")
    so.write_capture(p, src, wav=wav)
    with open(p, "ab") as f:                                   # a ragged tail: less than one block, dropped on replay
        f.write(b"\x01" * 1000)
    fs = FileSource(p) if wav else FileSource(p, RATE, fmt)
    assert (fs.rate, fs.fmt, fs.block_len, fs.kind) == (RATE, fmt, src.block_len, so.KIND_IQ)
    got = list(fs)
    assert len(got) == fs.n_blocks == src.n_blocks
    for b, blk in enumerate(got):
        want = src.block(b)
        assert np.array_equal(blk.view(want.dtype) if fmt == FMT_F32 else blk, want)


def test_wav_header_is_what_airspy_rx_writes(tmp_path):
    """t_wav_file_hdr (airspy_rx.c:126-159): 44 bytes, fmt chunk of 16, sizes patched at exit (:1084-1100)."""
    h = so.wav_header(1, 2, 10000000, 16, 4000)
    assert len(h) == so.WAV_HEADER_LEN == 44
    assert h[:4] == b"RIFF" and h[8:16] == b"WAVEfmt " and h[36:40] == b"data"
    riff, = struct.unpack("<I", h[4:8])
    clen, tag, ch, rate, bps, align, bits = struct.unpack("<IHHIIHH", h[16:36])
    dlen, = struct.unpack("<I", h[40:44])
    assert (riff, clen, tag, ch, rate, bps, align, bits, dlen) == (4036, 16, 1, 2, 10000000, 40000000, 4, 16, 4000)
    p = str(tmp_path / "a.wav")
    open(p, "wb").write(h + b"\0" * 4000)
    w = so.read_wav_header(p)
    assert (w.format_tag, w.channels, w.rate, w.bits, w.data_bytes, w.data_offset) == (1, 2, 10000000, 16, 4000, 44)


def test_wav_with_extra_chunks_and_unpatched_sizes(tmp_path):
    """SDR#-style files carry chunks between fmt and data; an interrupted airspy_rx leaves the sizes 0."""
    src = band(FMT_S16, 2)
    body = b"".join(np.ascontiguousarray(b).tobytes() for b in src.blocks())
    fmt_ck = b"fmt " + struct.pack("<IHHIIHH", 16, 1, 2, RATE, RATE * 4, 4, 16)
    extra = b"auxi" + struct.pack("<I", 5) + b"hello\0"                       # odd length: padded to even
    p = str(tmp_path / "b.wav")
    open(p, "wb").write(b"RIFF" + struct.pack("<I", 0) + b"WAVE" + fmt_ck + extra + b"data" + struct.pack("<I", 0) + body)
    fs = FileSource(p)
    assert fs.n_blocks == 2 and fs.fmt == FMT_S16 and fs.offset == 12 + 24 + 14 + 8
    assert all(np.array_equal(a, src.block(i)) for i, a in enumerate(fs))


def test_raw_capture_needs_rate_and_format(tmp_path):
    p = str(tmp_path / "c.bin")
    open(p, "wb").write(b"\0" * 100)
    with pytest.raises(ValueError):
        FileSource(p)
    with pytest.raises(ValueError):
        FileSource(p, RATE)
    assert FileSource(p, RATE, FMT_U8).n_blocks == 0


def test_sample_conventions():
    """u8: (x - 127.5) / 127.5 (rtl_dev.cpp:320-323); s16: x / 32768 (airspy.c:58-62): quantize() is their inverse to
    within half a step."""
    x = band(FMT_F32, 1).block_complex(0)[:4096]
    ri = np.stack([x.real, x.imag], axis=1).ravel().astype(np.float64)
    for fmt, conv, step in ((FMT_U8, lambda q: (q - 127.5) / 127.5, 1 / 127.5), (FMT_S16, lambda q: q / 32768.0, 1 / 32768.0)):
        q = so.quantize(x, fmt)
        assert q.dtype == (np.uint8 if fmt == FMT_U8 else np.int16) and q.size == 2 * x.size
        assert np.max(np.abs(conv(q.astype(np.float64)) - ri)) <= 0.5 * step + 1e-7


@pytest.mark.parametrize("kind", [so.KIND_REAL_U16, so.KIND_REAL_S16, so.KIND_REAL_PACKED12])
def test_real_sample_captures_feed_the_front_end(tmp_path, kind):
    """airspy_rx -t 3|4|5: one block of a real capture is 2 * block_len ADC samples; INT16_REAL is (adc-2048)<<4."""
    from tests.test_airspy_fe import adc_stream
    rate, L = 10000000, 320000
    x = adc_stream(2 * L * 2 + 1000, seed=4)
    p = str(tmp_path / "real.bin")
    if kind == so.KIND_REAL_U16:
        x.tofile(p)
    elif kind == so.KIND_REAL_S16:
        ((x.astype(np.int32) - 2048) << 4).astype(np.int16).tofile(p)
    else:
        ob.airspy_pack12(x[: x.size // 8 * 8]).tofile(p)
    fs = FileSource(p, rate, kind=kind)
    assert fs.n_blocks == 2 and fs.packed == (kind == so.KIND_REAL_PACKED12)
    for b, blk in enumerate(fs):
        want = x[2 * L * b: 2 * L * (b + 1)]
        if fs.packed:
            blk = ob.OracleAirspyFE(np.ones(47, np.float32)).unpack12(blk)
        assert blk.dtype == np.uint16 and np.array_equal(blk, want)


def test_real_wav_capture_carries_the_adc_rate(tmp_path):
    from tests.test_airspy_fe import adc_stream
    x = adc_stream(2 * 320000, seed=2)
    p = str(tmp_path / "real.wav")
    open(p, "wb").write(so.wav_header(1, 1, 20000000, 16, x.nbytes) + x.tobytes())
    fs = FileSource(p)
    assert (fs.kind, fs.rate, fs.n_blocks) == (so.KIND_REAL_U16, 10000000, 1)
    assert np.array_equal(next(iter(fs)), x)


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("fmt,wav", [(FMT_U8, False), (FMT_S16, True), (FMT_F32, True)], ids=["rtl_u8", "s16_wav", "f32_wav"])
def test_replayed_capture_matches_the_oracle(tmp_path, fmt, wav):
    from sdrx_b200 import ChannelPlan, Engine
    from tests.harness import AUDIO_TOL, product_tables, rel_rms
    offs = [-3, 0, 7]
    src = SyntheticSource(RATE, [Carrier(-3, 0.4, 1000.0, 0.5, lambda b: 1 <= b < 5), Carrier(7, 0.1, 600.0, 0.3)],
                          0.01, 99, fmt, 6)
    p = str(tmp_path / "cap")
    so.write_capture(p, src, wav=wav)
    fs = FileSource(p) if wav else FileSource(p, RATE, fmt)
    orx = ob.OracleRx(product_tables(RATE), offs)
    eng = Engine(ChannelPlan.from_offsets(RATE, offs))
    opened = False
    for blk in fs:
        want = orx.process(blk, fs.fmt)
        got = eng.process(blk, fs.fmt)
        assert np.array_equal(got.sql, want.sql)
        opened |= bool(want.sql.any())
        assert rel_rms(got.ch_iq, want.ch_iq) < AUDIO_TOL
        if np.any(want.mix):
            assert rel_rms(got.mix, want.mix) < AUDIO_TOL
    assert opened
    eng.close()


@pytest.mark.gpu
def test_replayed_packed_real_capture_through_the_front_end(tmp_path):
    """airspy_rx -t 5 -p 1 capture (12-bit packed ADC words) -> FileSource -> AirspyFrontEnd -> Engine == oracle chain."""
    from sdrx_b200 import AirspyFrontEnd, ChannelPlan, Engine
    from tests.harness import AUDIO_TOL, product_tables, rel_rms
    from tests.test_airspy_fe import hb
    rate, L, offs = 10000000, 320000, [-200, 5, 333]
    rng = np.random.Generator(np.random.PCG64(8))
    t = np.arange(2 * L * 3, dtype=np.float64)
    x = rng.normal(0.0, 5.0, t.size) + 2048 - 13
    for k, off in enumerate(offs):
        x += (80.0 + 30 * k) * (1 + 0.4 * np.cos(2 * np.pi * (500.0 + 400 * k) / 20e6 * t)) * \
            np.cos(2 * np.pi * (0.25 - off * (25000.0 / 3) / 20e6) * t)
    x = np.clip(np.rint(x), 0, 4095).astype(np.uint16)
    p = str(tmp_path / "packed.bin")
    ob.airspy_pack12(x).tofile(p)
    fs = FileSource(p, rate, kind=so.KIND_REAL_PACKED12)
    assert fs.n_blocks == 3
    ofe, orx = ob.OracleAirspyFE(hb()), ob.OracleRx(product_tables(rate), offs)
    fe, eng = AirspyFrontEnd(max_samples=2 * L, packed=True), Engine(ChannelPlan.from_offsets(rate, offs))
    opened = False
    for blk in fs:
        want = orx.process(ofe.process(ofe.unpack12(blk)), 2)
        fe.feed(eng, blk)
        got = eng.outputs()
        assert np.array_equal(got.sql, want.sql)
        opened |= bool(want.sql.any())
        assert rel_rms(eng.channel_iq(), want.ch_iq) < AUDIO_TOL
    assert opened
    fe.close()
    eng.close()


def test_replay_tool_arguments():
    """tools/replay.py takes the channel syntax of the sdrx command line and the capture kinds of FileSource."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("replay", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "replay.py"))
    rp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rp)
    a = rp.build_parser().parse_args(["cap.bin", "118.105/12/FM", "121.500", "--rate", "10000000", "--real", "packed12", "-o", "x.wav"])
    assert a.channels == ["118.105/12/FM", "121.500"] and a.real == "packed12" and a.rate == 10000000
    h = rp.pcm_wav_header(512)
    assert len(h) == 44 and struct.unpack("<IHHIIHH", h[16:36]) == (16, 1, 2, 16000, 64000, 4, 16)
    assert struct.unpack("<I", h[40:44])[0] == 2048


@pytest.mark.gpu
def test_replay_tool_end_to_end(tmp_path, capsys):
    """A synthetic rtl_sdr-style capture replayed by tools/replay.py: the WAV it writes is the engine's pcm stream."""
    import importlib.util
    from sdrx_b200 import ChannelPlan, Engine
    spec = importlib.util.spec_from_file_location("replay", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "replay.py"))
    rp = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rp)
    src = SyntheticSource(RATE, [Carrier(-3, 0.4, 1000.0, 0.5, lambda b: 1 <= b < 6)], 0.01, 7, FMT_U8, 8)
    cap, wav = str(tmp_path / "rtl.bin"), str(tmp_path / "audio.wav")
    so.write_capture(cap, src)
    assert rp.main([cap, "118.275", "--rate", str(RATE), "--format", "u8", "-o", wav]) == 0
    text = capsys.readouterr().out
    assert "open" in text and "close" in text and "118.275" in text
    eng = Engine(ChannelPlan.from_names(RATE, ["118.275"]))
    want = np.concatenate([eng.process(src.block(b), FMT_U8).pcm for b in range(8)])
    eng.close()
    w = so.read_wav_header(wav)
    assert (w.format_tag, w.channels, w.rate, w.bits, w.data_bytes) == (1, 2, 16000, 16, want.size * 2)
    got = np.fromfile(wav, dtype="<i2", offset=w.data_offset)
    assert np.array_equal(got, want) and np.any(got != 0)

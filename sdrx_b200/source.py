"""Deterministic IQ sources that stand in for the dongle (r820_dev / rtl_dev / airspy_dev).

The reference's device layer emits one block of ``512 * M`` complex samples every 32 ms through
``sigc::signal<void(const iqsample_t*, unsigned, void*, const BlockInfo&)> data``
(/root/reference/src/r820_dev.hpp:134-139).  :class:`SyntheticSource` and :class:`FileSource` deliver
the same blocks -- in the dongle's *wire* format (uint8 / int16 / float32 IQ), because the B200
engine fuses the raw->float step -- either by iteration or through :meth:`connect` + :meth:`run`
with the reference's callback argument list.  A zero-length block with ``stream_state == IDLE``
marks the end of the stream (sdrx.cpp:313-317).
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Callable, Iterator, List, Optional, Sequence

import numpy as np

from ._lib import FMT_F32, FMT_S16, FMT_U8
from .config import rate_config

STREAMING, IDLE = 1, 0      # R820Dev::StreamState (r820_dev.hpp:76-79)


@dataclass
class BlockInfo:            # r820_dev.hpp:81-97
    stream_state: int
    rate: int
    pwr: float
    ts: float


@dataclass
class Carrier:
    """One AM transmitter on the 8.33 kHz grid."""
    offset: int                         # channel offset in 8.33 kHz steps from the tuner centre
    amp: float = 0.1                    # carrier amplitude (full scale = 1)
    tone_hz: float = 1000.0             # modulating tone
    depth: float = 0.5                  # AM depth
    on: Optional[Callable[[int], bool]] = None   # block index -> keyed?  None = always on
    df_hz: float = 0.0                  # carrier offset inside the channel
    fm_dev_hz: float = 0.0              # FM: peak deviation of the tone (0 = none); use depth=0 for pure FM


def quantize(x: np.ndarray, fmt: int) -> np.ndarray:
    """complex64 -> wire format (inverse of rtl_dev.cpp:320-323 / the airspy int16 convention)."""
    if fmt == FMT_F32:
        return np.ascontiguousarray(x, dtype=np.complex64)
    iq = np.empty(2 * x.size, dtype=np.float32)
    iq[0::2], iq[1::2] = x.real, x.imag
    if fmt == FMT_U8:
        return np.clip(np.rint((iq + 1.0) * 127.5), 0, 255).astype(np.uint8)
    return np.clip(np.rint(iq * 32768.0), -32768, 32767).astype(np.int16)


class _SourceBase:
    rate: int
    block_len: int
    fmt: int

    def __init__(self):
        self._cb = None
        self._user = None

    def blocks(self) -> Iterator[np.ndarray]:
        raise NotImplementedError

    def __iter__(self):
        return self.blocks()

    def connect(self, cb: Callable, user_data=None) -> None:
        """`device->data.connect(sigc::ptr_fun(data_cb))` (sdrx.cpp:1772)."""
        self._cb, self._user = cb, user_data

    def run(self) -> None:
        """`device->start()`: emit every block, then the zero-length IDLE block."""
        ts = 0.0
        for blk in self.blocks():
            n = blk.size if blk.dtype == np.complex64 else blk.size // 2
            self._cb(blk, n, self._user, BlockInfo(STREAMING, self.rate, float("nan"), ts))
            ts += 0.032
        self._cb(None, 0, self._user, BlockInfo(IDLE, self.rate, 0.0, ts))


class SyntheticSource(_SourceBase):
    """Seeded synthetic band: AM carriers on chosen channels plus complex white noise.

    Block ``b`` is a pure function of ``(seed, b)`` so any block can be regenerated on any rank."""

    def __init__(self, rate: int, carriers: Sequence[Carrier], noise_sigma: float = 0.01, seed: int = 12345,
                 fmt: int = FMT_F32, n_blocks: int = 8, block_len: Optional[int] = None):
        super().__init__()
        self.rate = rate
        self.cfg = rate_config(rate)
        self.block_len = int(block_len or self.cfg.block_len)
        self.carriers = list(carriers)
        self.noise_sigma = float(noise_sigma)
        self.seed = seed
        self.fmt = fmt
        self.n_blocks = n_blocks

    def block_complex(self, b: int) -> np.ndarray:
        L, N = self.block_len, self.cfg.N
        n0 = b * L
        n = np.arange(n0, n0 + L, dtype=np.int64)
        t = n.astype(np.float64) / float(self.rate)
        x = np.zeros(L, dtype=np.complex128)
        step_hz = self.rate * self.cfg.z / N          # 8333.33 Hz (sdrx.cpp:1565-1569)
        for c in self.carriers:
            if c.on is not None and not c.on(b):
                continue
            # grid part of the phase kept exact with integers; small df added in float
            ph = 2.0 * np.pi * (((c.offset * self.cfg.z) * n) % N).astype(np.float64) / N
            if c.df_hz:
                ph = ph + 2.0 * np.pi * c.df_hz * t
            if c.fm_dev_hz:             # phase = integral of 2 pi dev cos(2 pi f t)
                ph = ph + (c.fm_dev_hz / c.tone_hz) * np.sin(2.0 * np.pi * c.tone_hz * t)
            env = c.amp * (1.0 + c.depth * np.sin(2.0 * np.pi * c.tone_hz * t))
            x += env * np.exp(1j * ph)
        _ = step_hz
        if self.noise_sigma > 0:
            rng = np.random.Generator(np.random.PCG64([self.seed, b]))
            x += self.noise_sigma * (rng.standard_normal(L) + 1j * rng.standard_normal(L))
        return x.astype(np.complex64)

    def block(self, b: int) -> np.ndarray:
        return quantize(self.block_complex(b), self.fmt)

    def blocks(self) -> Iterator[np.ndarray]:
        for b in range(self.n_blocks):
            yield self.block(b)


WAV_HEADER_LEN = 44      # t_wav_file_hdr: RIFF(12) + 'fmt '(24) + 'data'(8)  (airspy_rx.c:126-159)

# Sample kinds of a capture.  The IQ kinds go straight to Engine.push (wire formats FMT_*); the REAL kinds are the
# ADC's 20 MS/s real stream and go through AirspyFrontEnd first (airspy.c:391-420, `airspy_rx -t 3|4|5`).
KIND_IQ, KIND_REAL_U16, KIND_REAL_S16, KIND_REAL_PACKED12 = "iq", "real_u16", "real_s16", "real_packed12"


@dataclass
class WavInfo:
    """The fields of airspy_rx's WAV header (airspy_rx.c:134-146, filled in at :1084-1096)."""
    format_tag: int         # 1 = PCM16, 3 = float32
    channels: int           # 2 = IQ, 1 = real
    rate: int
    bits: int
    data_bytes: int
    data_offset: int


def read_wav_header(path: str) -> Optional[WavInfo]:
    """Parses a RIFF/WAVE header (any chunk order, as SDR# writes extra chunks); None when the file is not RIFF."""
    import struct
    with open(path, "rb") as f:
        head = f.read(12)
        if len(head) < 12 or head[:4] != b"RIFF" or head[8:12] != b"WAVE":
            return None
        fmt = None
        size = os.path.getsize(path)
        while True:
            ck = f.read(8)
            if len(ck) < 8:
                raise ValueError(f"{path}: WAV file without a data chunk")
            cid, clen = ck[:4], struct.unpack("<I", ck[4:])[0]
            if cid == b"fmt ":
                body = f.read(clen + (clen & 1))
                fmt = struct.unpack("<HHIIHH", body[:16])
            elif cid == b"data":
                if fmt is None:
                    raise ValueError(f"{path}: WAV data chunk before the fmt chunk")
                off = f.tell()
                # a capture that was interrupted keeps chunkSize 0 (the header is patched on exit): trust the file
                n = clen if 0 < clen <= size - off else size - off
                return WavInfo(fmt[0], fmt[1], fmt[2], fmt[5], n, off)
            else:
                f.seek(clen + (clen & 1), os.SEEK_CUR)


def wav_header(format_tag: int, channels: int, rate: int, bits: int, data_bytes: int) -> bytes:
    """The 44-byte header airspy_rx writes (airspy_rx.c:161-190, :1080-1100)."""
    import struct
    align = channels * bits // 8
    return (b"RIFF" + struct.pack("<I", (WAV_HEADER_LEN + data_bytes - 8) & 0xffffffff) + b"WAVE" + b"fmt "
            + struct.pack("<IHHIIHH", 16, format_tag, channels, rate, (rate * align) & 0xffffffff, align, bits)
            + b"data" + struct.pack("<I", data_bytes & 0xffffffff))


class FileSource(_SourceBase):
    """Replay of a capture: rtl_sdr raw u8 IQ (librtlsdr/src/rtl_sdr.c:113,:329); airspy_rx raw or WAV in INT16_IQ
    (its default, airspy_rx.c:102) / FLOAT32_IQ (airspy_rx.c:374-387, :592-617); and airspy_rx's REAL sample types
    (UINT16_REAL, INT16_REAL, RAW 12-bit packed: airspy.c:409-420, airspy_rx.c:389-409), whose blocks are ADC samples
    at twice the IQ rate for :class:`AirspyFrontEnd`.  A WAV header is detected and skipped, and supplies format and
    rate unless given.  A trailing partial block is dropped, as the device layer only ever emits whole 32 ms blocks."""

    def __init__(self, path: str, rate: Optional[int] = None, fmt: Optional[int] = None, block_len: Optional[int] = None,
                 kind: Optional[str] = None):
        super().__init__()
        self.path = path
        self.wav = read_wav_header(path)
        self.offset = 0
        if self.wav is not None:
            w = self.wav
            self.offset = w.data_offset
            if kind is None:
                kind = KIND_IQ if w.channels == 2 else KIND_REAL_U16
            if fmt is None and kind == KIND_IQ:
                try:
                    fmt = {(1, 16): FMT_S16, (3, 32): FMT_F32, (1, 8): FMT_U8}[(w.format_tag, w.bits)]
                except KeyError:
                    raise ValueError(f"{path}: unsupported WAV sample format tag {w.format_tag}, {w.bits} bits")
            if rate is None:
                rate = w.rate if w.channels == 2 else w.rate // 2     # a real capture carries the ADC rate
        if rate is None:
            raise ValueError("a raw capture carries no rate: pass rate=")
        self.kind = kind or KIND_IQ
        self.rate = int(rate)
        self.block_len = int(block_len or rate_config(self.rate).block_len)
        if self.kind == KIND_IQ:
            if fmt is None:
                raise ValueError("a raw IQ capture carries no sample format: pass fmt=")
            self.fmt = fmt
            self.dtype = {FMT_U8: np.uint8, FMT_S16: np.int16, FMT_F32: np.float32}[fmt]
            self.items_per_block = self.block_len * 2
        else:
            self.fmt = FMT_F32                                   # what the front end hands the engine
            self.dtype = {KIND_REAL_U16: np.uint16, KIND_REAL_S16: np.int16, KIND_REAL_PACKED12: np.uint32}[self.kind]
            n_adc = self.block_len * 2                           # real samples per block
            self.items_per_block = n_adc * 3 // 8 if self.kind == KIND_REAL_PACKED12 else n_adc
        size = (self.wav.data_bytes if self.wav is not None else os.path.getsize(path) - self.offset)
        self.n_blocks = size // (self.items_per_block * np.dtype(self.dtype).itemsize)

    @property
    def packed(self) -> bool:
        return self.kind == KIND_REAL_PACKED12

    def blocks(self) -> Iterator[np.ndarray]:
        with open(self.path, "rb") as f:
            f.seek(self.offset)
            for _ in range(self.n_blocks):
                blk = np.fromfile(f, dtype=self.dtype, count=self.items_per_block)
                if self.kind == KIND_REAL_S16:
                    # INT16_REAL is (adc - 2048) << 4 (airspy.c:58-61, :299-309, SAMPLE_SHIFT 4): undo it, exactly
                    blk = ((blk >> 4) + 2048).astype(np.uint16)
                yield blk


def write_capture(path: str, src: _SourceBase, wav: bool = False) -> None:
    """Writes ``src``'s blocks the way rtl_sdr / airspy_rx would: raw, or (``wav``) behind airspy_rx's WAV header."""
    blocks = [np.ascontiguousarray(b) for b in src.blocks()]
    with open(path, "wb") as f:
        if wav:
            tag, bits = {FMT_S16: (1, 16), FMT_F32: (3, 32), FMT_U8: (1, 8)}[src.fmt]
            f.write(wav_header(tag, 2, src.rate, bits, sum(b.nbytes for b in blocks)))
        for blk in blocks:
            blk.tofile(f)

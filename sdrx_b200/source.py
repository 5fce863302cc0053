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


class FileSource(_SourceBase):
    """Replay of a raw capture: rtl_sdr u8 IQ (librtlsdr/src/rtl_sdr.c:113,:329), airspy_rx INT16_IQ /
    FLOAT32_IQ (libairspy/airspy-tools/src/airspy_rx.c:481-482).  A trailing partial block is dropped,
    as the device layer only ever emits whole 32 ms blocks."""

    def __init__(self, path: str, rate: int, fmt: int, block_len: Optional[int] = None):
        super().__init__()
        self.path, self.rate, self.fmt = path, rate, fmt
        self.block_len = int(block_len or rate_config(rate).block_len)
        self.dtype = {FMT_U8: np.uint8, FMT_S16: np.int16, FMT_F32: np.float32}[fmt]
        self.n_blocks = os.path.getsize(path) // (self.block_len * 2 * np.dtype(self.dtype).itemsize)

    def blocks(self) -> Iterator[np.ndarray]:
        with open(self.path, "rb") as f:
            for _ in range(self.n_blocks):
                yield np.fromfile(f, dtype=self.dtype, count=self.block_len * 2)


def write_capture(path: str, src: SyntheticSource) -> None:
    with open(path, "wb") as f:
        for blk in src.blocks():
            np.ascontiguousarray(blk).tofile(f)

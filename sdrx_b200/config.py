"""Host-side set-up logic of sdrx, mirrored for the B200 engine (reference: /root/reference/src/sdrx.cpp).

* :func:`rate_config`      -- N, z and the MSD stage list per sample rate (sdrx.cpp:1584-1677)
* :func:`parse_fq`         -- aeronautical designator / frequency parser (sdrx.cpp:962-1019)
* :func:`channel_to_offset`-- designator -> 8.33 kHz steps from the tuner (sdrx.cpp:1449-1478)
* :func:`tuner_fq_for`     -- tuner centre: midpoint of lowest/highest channel, rounded to 100 kHz (:1398-1408)
* :func:`get_audio_pos`    -- pan position per channel index (sdrx.cpp:1024-1044)
* :func:`make_translator`  -- the translator table (sdrx.cpp:1682-1690), evaluated by the C library with
  the host's libm so that it matches the reference bit for bit.

The coefficient tables are data extracted from the reference headers by tools/extract_tables.py
into sdrx_b200/data/tables.npz.
"""
from __future__ import annotations

import math
import os
import re
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "tables.npz")
_TABLES = None

SAMPLE_RATES = (960000, 1200000, 1440000, 1600000, 1920000, 2400000, 2560000, 6000000, 10000000)
CH_IQ_BUF_SIZE = 512            # sdrx.cpp:129
CH_IQ_SAMPLING_FQ = 16000       # rates.hpp:31-44: every chain ends at 16 kS/s


def tables():
    global _TABLES
    if _TABLES is None:
        with np.load(_DATA) as z:
            _TABLES = {k: z[k] for k in z.files}
    return _TABLES


@dataclass
class RateConfig:
    rate: int
    N: int                      # translator length
    z: int
    stages: List[Tuple[int, np.ndarray]]   # [(m, taps)]

    @property
    def M(self) -> int:         # total decimation, rate / 16000
        out = 1
        for m, _ in self.stages:
            out *= m
        return out

    @property
    def block_len(self) -> int:  # complex samples per 32 ms block (rtl_dev.cpp:204-211)
        return CH_IQ_BUF_SIZE * self.M


def rate_config(rate: int) -> RateConfig:
    t = tables()
    if f"rate_{rate}_N" not in t:
        raise ValueError(f"Sample rate {rate} is not supported (sdrx.cpp:1674-1677)")
    m = t[f"rate_{rate}_m"]
    nt = t[f"rate_{rate}_ntaps"]
    taps = t[f"rate_{rate}_taps"]
    stages, pos = [], 0
    for s in range(len(m)):
        stages.append((int(m[s]), taps[pos:pos + int(nt[s])].astype(np.float32)))
        pos += int(nt[s])
    return RateConfig(rate, int(t[f"rate_{rate}_N"]), int(t[f"rate_{rate}_z"]), stages)


# sdrx.cpp:991-996 (Hz) and :1464-1469 (8.33 kHz steps)
_SUB_HZ = {"00": 0, "05": 0, "10": 8333, "15": 16667, "25": 25000, "30": 25000, "35": 33333, "40": 41667,
           "50": 50000, "55": 50000, "60": 58333, "65": 66667, "75": 75000, "80": 75000, "85": 83333, "90": 91667}
_SUB_STEP = {"00": 0, "05": 0, "10": 1, "15": 2, "25": 3, "30": 3, "35": 4, "40": 5,
             "50": 6, "55": 6, "60": 7, "65": 8, "75": 9, "80": 9, "85": 10, "90": 11}
_STEP_SUB = {0: "05", 1: "10", 2: "15", 3: "30", 4: "35", 5: "40", 6: "55", 7: "60", 8: "65", 9: "80", 10: "85", 11: "90"}


def parse_fq(s: str, aeronautical: bool = False) -> int:
    """sdrx.cpp:962-1019.  Returns Hz, or 0 for an invalid string."""
    if "." not in s:
        return 0
    int_str, frac_str = s.split(".", 1)
    if (not int_str.isdigit() or not frac_str.isdigit() or not (2 <= len(int_str) <= 4)
            or not (1 <= len(frac_str) <= 6) or not int_str.isascii() or not frac_str.isascii()):
        return 0
    if aeronautical and len(frac_str) != 3:
        return 0
    mhz = hz = 0
    if aeronautical:
        sub = _SUB_HZ.get(frac_str[1:])
        if sub is not None:
            mhz = int(int_str)
            hz = int(frac_str[0]) * 100000 + sub
    else:
        mhz = int(int_str)
        for d, mult in zip(frac_str, (100000, 10000, 1000, 100, 10, 1)):
            hz += int(d) * mult
    return mhz * 1000000 + hz if mhz < 4000 else 0


def channel_to_offset(channel: str, tuner_fq: int) -> int:
    """sdrx.cpp:1449-1478 (C integer division truncates toward zero)."""
    int_str, frac_str = channel.split(".", 1)
    fq_base = int(int_str) * 1000000 + int(frac_str[0]) * 100000
    fq_diff = fq_base - int(tuner_fq)
    q = abs(fq_diff) // 100000
    offset_diff = (-q if fq_diff < 0 else q) * 12
    return offset_diff + _SUB_STEP[frac_str[1:]]


def offset_to_channel(offset: int, tuner_fq: int) -> str:
    """Inverse of :func:`channel_to_offset` on the 8.33 kHz grid (tuner_fq a multiple of 100 kHz)."""
    base, step = divmod(offset, 12)
    hz100 = tuner_fq // 100000 + base
    return f"{hz100 // 10:03d}.{hz100 % 10}{_STEP_SUB[step]}"


def tuner_fq_for(channels: Sequence[str]) -> int:
    """sdrx.cpp:1398-1408."""
    names = sorted(channels)
    lo, hi = float(parse_fq(names[0], True)), float(parse_fq(names[-1], True))
    mid = lo + (hi - lo) / 2
    return int(_c_round(mid / 100000.0) * 100000.0)


def _c_round(x: float) -> float:     # std::round: half away from zero
    return math.floor(x + 0.5) if x >= 0 else math.ceil(x - 0.5)


def get_audio_pos(channel_no: int, num_channels: int) -> int:
    """sdrx.cpp:1024-1044."""
    num_positions = 5
    half = num_channels // 2
    odd = num_channels % 2
    pos = 0
    if channel_no < num_channels:
        if channel_no < half:
            tmp = np.float32(channel_no * num_positions) / np.float32(num_channels)
            pos = int(math.floor(tmp)) - num_positions // 2
        elif channel_no == half and odd:
            pos = 0
        else:
            tmp = np.float32((num_channels - 1 - channel_no) * num_positions) / np.float32(num_channels)
            pos = num_positions // 2 - int(math.floor(tmp))
    return pos


_CH_RE = re.compile(r"^([0-9]{3}\.[0-9]{3})(?:/([0-9]{1,2})(?:/(AM|FM))?)?$")   # sdrx.cpp:1351


def parse_channel_arg(arg: str, default_sql: float = 9.0):
    """`NNN.NNN[/SQL[/AM|FM]]` -> (name, sql_level, modulation) or None (sdrx.cpp:1349-1382)."""
    mt = _CH_RE.match(arg)
    if not mt:
        return None
    name = mt.group(1)
    sql = float(mt.group(2)) if mt.group(2) else default_sql
    mod = mt.group(3) or "AM"
    fq = parse_fq(name, True)
    if fq == 0 or fq < 45000000 or fq > 1800000000 or sql < 0.0 or sql > 50.0:
        return None
    return name, sql, mod


def make_translator(N: int, z: int, ch_offset: int) -> np.ndarray:
    """sdrx.cpp:1682-1690 via the C library (same libm as the reference).  Empty for offset 0."""
    if ch_offset == 0:
        return np.zeros(0, dtype=np.complex64)
    from . import _lib
    out = np.zeros(N, dtype=np.complex64)
    _lib.load().sdrx_make_translator(N, z, int(ch_offset), out.view(np.float32).ctypes.data_as(_lib.c_float_p))
    return out


def make_window() -> np.ndarray:
    from . import _lib
    out = np.zeros(CH_IQ_BUF_SIZE, dtype=np.float32)
    _lib.load().sdrx_make_window(out.ctypes.data_as(_lib.c_float_p))
    return out


@dataclass
class ChannelPlan:
    """What sdrx.cpp:1680-1714 derives for a list of channels."""
    rate: RateConfig
    tuner_fq: int
    names: List[str]
    offsets: List[int]
    sql_levels: List[float]
    pos: List[int] = field(default_factory=list)
    modulation: List[str] = field(default_factory=list)     # "AM" / "FM" per channel (Channel::mod, sdrx.cpp:230); [] -> AM

    @classmethod
    def from_names(cls, rate: int, args: Sequence[str], default_sql: float = 9.0, default_mod: str = "AM",
                   tuner_fq: Optional[int] = None) -> "ChannelPlan":
        """``default_mod`` is --modulation (sdrx.cpp:1197, :1358); a channel's own /AM|/FM suffix wins.  ``tuner_fq``
        overrides the centre frequency sdrx would pick for these channels (a capture was made at a fixed one)."""
        parsed = []
        for a in args:
            p = parse_channel_arg(a, default_sql)
            if p is None:
                raise ValueError(f"Invalid channel given: {a}")
            if not _CH_RE.match(a).group(3):
                p = (p[0], p[1], default_mod)
            if p[0] not in [q[0] for q in parsed]:       # sdrx.cpp:1377-1380 de-duplicates
                parsed.append(p)
        names = [p[0] for p in parsed]
        tuner = tuner_fq_for(names) if tuner_fq is None else int(tuner_fq)
        offs = [channel_to_offset(n, tuner) for n in names]
        return cls(rate_config(rate), tuner, names, offs, [p[1] for p in parsed],
                   [get_audio_pos(i, len(names)) for i in range(len(names))], [p[2] for p in parsed])

    @classmethod
    def from_offsets(cls, rate: int, offsets: Sequence[int], sql_levels: Optional[Sequence[float]] = None,
                     tuner_fq: int = 122000000, modulation: Optional[Sequence[str]] = None) -> "ChannelPlan":
        offs = [int(o) for o in offsets]
        return cls(rate_config(rate), tuner_fq, [offset_to_channel(o, tuner_fq) for o in offs], offs,
                   list(sql_levels) if sql_levels is not None else [9.0] * len(offs),
                   [get_audio_pos(i, len(offs)) for i in range(len(offs))],
                   list(modulation) if modulation is not None else [])

    def modulation_codes(self) -> List[int]:
        """0 AM / 1 FM per channel (SDRX_MOD_*)."""
        mods = self.modulation or ["AM"] * len(self.offsets)
        if len(mods) != len(self.offsets) or any(m not in ("AM", "FM") for m in mods):
            raise ValueError("modulation must be 'AM' or 'FM' for every channel")
        return [1 if m == "FM" else 0 for m in mods]

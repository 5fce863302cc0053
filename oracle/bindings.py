"""TEST INFRASTRUCTURE -- ctypes bindings for the CPU oracle libraries.

``oracle/liboracle.so``            plain-C restatement (sdrx_oracle.c, fft512.c)       backend "oracle"
``oracle/_ref/libsdrx_ref.so``     the reference's own headers, pinned portable flags   backend "ref"
``oracle/_ref/libsdrx_ref_avx2.so`` same harness, the reference's Release flags (AVX2)   backend "ref_avx2"

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  Nothing under ``sdrx_b200/`` does.

All complex data cross the boundary as ``np.complex64`` (== ``iqsample_t``,
/root/reference/src/iqsample.hpp:26).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"

FMT_U8, FMT_S16, FMT_F32 = 0, 1, 2
CH_IQ_BUF_SIZE = 512

_PATHS = {
    "oracle": os.path.join(HERE, "liboracle.so"),
    "ref": os.path.join(HERE, "_ref", "libsdrx_ref.so"),
    "ref_avx2": os.path.join(HERE, "_ref", "libsdrx_ref_avx2.so"),
}

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_u8_p = C.POINTER(C.c_uint8)
c_i16_p = C.POINTER(C.c_int16)


def build(force: bool = False) -> None:
    """Compile liboracle.so and, where /root/reference exists, oracle/_ref/*.so (oracle/Makefile)."""
    args = ["make", "-C", HERE, "-s"]
    if force:
        args.append("-B")
    subprocess.run(args, check=True)


def backend_available(name: str) -> bool:
    return os.path.exists(_PATHS[name])


def _fp(a: np.ndarray):
    return a.ctypes.data_as(c_float_p)


@dataclass
class Backend:
    name: str
    lib: C.CDLL

    @property
    def is_ref(self) -> bool:
        return self.name.startswith("ref")


_CACHE: Dict[str, Backend] = {}


def load_backend(name: str = "oracle") -> Backend:
    if name in _CACHE:
        return _CACHE[name]
    path = _PATHS[name]
    if not os.path.exists(path):
        if name == "oracle" or os.path.isdir(REF_SRC):
            build()
    if not os.path.exists(path):
        raise FileNotFoundError(f"oracle backend {name!r}: {path} is missing (run `make -C oracle`)")
    lib = C.CDLL(path)
    if name == "oracle":
        _declare_oracle(lib)
    else:
        _declare_ref(lib)
    be = Backend(name, lib)
    _CACHE[name] = be
    return be


# ----------------------------------------------------------------------------- prototypes

class _AgcS(C.Structure):
    _fields_ = [(n, C.c_float) for n in
                ("attack", "decay", "reference", "min_gain", "max_gain", "alpha", "gain", "power")]


class _LfAgcS(C.Structure):
    _fields_ = [(n, C.c_float) for n in
                ("attack", "decay", "reference", "min_gain", "max_gain", "alpha", "gain", "power")] + \
               [("active", C.c_int)]


class _RxConfig(C.Structure):
    _fields_ = [
        ("n_channels", C.c_int), ("N", C.c_int),
        ("translators", c_float_p), ("has_translator", c_u8_p),
        ("n_stages", C.c_int), ("m", c_int_p), ("ntaps", c_int_p), ("taps", c_float_p),
        ("sql_level", c_float_p), ("pos", c_int_p),
        ("use_lf_agc", C.c_int), ("use_ftfir", C.c_int), ("lf_gain_db", C.c_float),
        ("fir2_taps", c_float_p), ("fir2_ntaps", C.c_int),
        ("passband_shape", c_float_p), ("ramp_up", c_float_p), ("ramp_down", c_float_p),
    ]


def _declare_oracle(lib: C.CDLL) -> None:
    lib.sdrxo_translator.argtypes = [C.c_int, C.c_int, C.c_int, c_float_p]
    lib.sdrxo_translator.restype = None
    lib.sdrxo_window.argtypes = [c_float_p]
    lib.sdrxo_window.restype = None
    lib.sdrxo_audio_pos.argtypes = [C.c_uint, C.c_uint]
    lib.sdrxo_audio_pos.restype = C.c_int
    lib.sdrxo_channel_to_offset.argtypes = [C.c_char_p, C.c_int32]
    lib.sdrxo_channel_to_offset.restype = C.c_int
    lib.sdrxo_parse_fq.argtypes = [C.c_char_p, C.c_int]
    lib.sdrxo_parse_fq.restype = C.c_uint32
    lib.sdrxo_msd_create.argtypes = [c_float_p, C.c_int, C.c_int, c_int_p, c_int_p, c_float_p, C.c_int]
    lib.sdrxo_msd_create.restype = C.c_void_p
    lib.sdrxo_msd_clone.argtypes = [C.c_void_p]
    lib.sdrxo_msd_clone.restype = C.c_void_p
    lib.sdrxo_msd_destroy.argtypes = [C.c_void_p]
    lib.sdrxo_msd_destroy.restype = None
    lib.sdrxo_msd_m.argtypes = [C.c_void_p]
    lib.sdrxo_msd_m.restype = C.c_size_t
    lib.sdrxo_msd_decimate.argtypes = [C.c_void_p, c_float_p, C.c_size_t, c_float_p]
    lib.sdrxo_msd_decimate.restype = C.c_size_t
    lib.sdrxo_agc_init.argtypes = [C.POINTER(_AgcS)] + [C.c_float] * 5
    lib.sdrxo_agc_init.restype = None
    lib.sdrxo_agc_set_max_gain.argtypes = [C.POINTER(_AgcS), C.c_float]
    lib.sdrxo_agc_set_max_gain.restype = None
    lib.sdrxo_agc_adjust.argtypes = [C.POINTER(_AgcS), c_float_p, C.c_uint, c_float_p]
    lib.sdrxo_agc_adjust.restype = None
    lib.sdrxo_lfagc_init.argtypes = [C.POINTER(_LfAgcS)] + [C.c_float] * 5
    lib.sdrxo_lfagc_init.restype = None
    lib.sdrxo_lfagc_set_max_gain.argtypes = [C.POINTER(_LfAgcS), C.c_float]
    lib.sdrxo_lfagc_set_max_gain.restype = None
    lib.sdrxo_lfagc_adjust.argtypes = [C.POINTER(_LfAgcS), c_float_p, C.c_uint, c_float_p]
    lib.sdrxo_lfagc_adjust.restype = None
    lib.sdrxo_fir_create.argtypes = [C.c_int, c_float_p, C.c_int]
    lib.sdrxo_fir_create.restype = C.c_void_p
    lib.sdrxo_fir_destroy.argtypes = [C.c_void_p]
    lib.sdrxo_fir_destroy.restype = None
    lib.sdrxo_fir_set_gain.argtypes = [C.c_void_p, C.c_float]
    lib.sdrxo_fir_set_gain.restype = None
    lib.sdrxo_fir_filter.argtypes = [C.c_void_p, c_float_p, C.c_uint, c_float_p]
    lib.sdrxo_fir_filter.restype = None
    lib.sdrxo_rx_create.argtypes = [C.POINTER(_RxConfig)]
    lib.sdrxo_rx_create.restype = C.c_void_p
    lib.sdrxo_rx_destroy.argtypes = [C.c_void_p]
    lib.sdrxo_rx_destroy.restype = None
    lib.sdrxo_rx_ingest.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint]
    lib.sdrxo_rx_ingest.restype = C.c_float
    lib.sdrxo_rx_channelize.argtypes = [C.c_void_p, c_float_p]
    lib.sdrxo_rx_channelize.restype = None
    lib.sdrxo_rx_demod.argtypes = [C.c_void_p, c_float_p, c_u8_p, c_float_p, c_float_p, c_float_p, c_i16_p]
    lib.sdrxo_rx_demod.restype = None
    lib.sdrxo_rx_gains.argtypes = [C.c_void_p, c_float_p, c_float_p]
    lib.sdrxo_rx_gains.restype = None
    lib.sdrxo_rx_set_modulation.argtypes = [C.c_void_p, c_int_p]
    lib.sdrxo_rx_set_modulation.restype = None
    lib.sdrx_oracle_fft512.argtypes = [c_float_p, c_float_p]
    lib.sdrx_oracle_fft512.restype = None


def _declare_ref(lib: C.CDLL) -> None:
    lib.sdrxref_rate_config.argtypes = [C.c_uint32, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p, c_float_p, C.c_int]
    lib.sdrxref_rate_config.restype = C.c_int
    lib.sdrxref_table.argtypes = [C.c_char_p, c_float_p, C.c_int]
    lib.sdrxref_table.restype = C.c_int
    lib.sdrxref_translator.argtypes = [C.c_int, C.c_int, C.c_int, c_float_p]
    lib.sdrxref_translator.restype = None
    lib.sdrxref_window.argtypes = [c_float_p]
    lib.sdrxref_window.restype = None
    lib.sdrxref_audio_pos.argtypes = [C.c_uint, C.c_uint]
    lib.sdrxref_audio_pos.restype = C.c_int
    lib.sdrxref_msd_create.argtypes = [c_float_p, C.c_int, C.c_int, c_int_p, c_int_p, c_float_p, C.c_int]
    lib.sdrxref_msd_create.restype = C.c_void_p
    lib.sdrxref_msd_clone.argtypes = [C.c_void_p]
    lib.sdrxref_msd_clone.restype = C.c_void_p
    lib.sdrxref_msd_destroy.argtypes = [C.c_void_p]
    lib.sdrxref_msd_destroy.restype = None
    lib.sdrxref_msd_m.argtypes = [C.c_void_p]
    lib.sdrxref_msd_m.restype = C.c_size_t
    lib.sdrxref_msd_decimate.argtypes = [C.c_void_p, c_float_p, C.c_size_t, c_float_p]
    lib.sdrxref_msd_decimate.restype = C.c_size_t
    lib.sdrxref_dsbank_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    lib.sdrxref_dsbank_create.restype = C.c_void_p
    lib.sdrxref_dsbank_destroy.argtypes = [C.c_void_p]
    lib.sdrxref_dsbank_destroy.restype = None
    lib.sdrxref_dsbank_run.argtypes = [C.c_void_p, c_float_p, C.c_uint, c_float_p]
    lib.sdrxref_dsbank_run.restype = None
    for kind in ("agc", "lfagc"):
        getattr(lib, f"sdrxref_{kind}_create").argtypes = [C.c_float] * 5
        getattr(lib, f"sdrxref_{kind}_create").restype = C.c_void_p
        getattr(lib, f"sdrxref_{kind}_destroy").argtypes = [C.c_void_p]
        getattr(lib, f"sdrxref_{kind}_destroy").restype = None
        getattr(lib, f"sdrxref_{kind}_gain").argtypes = [C.c_void_p]
        getattr(lib, f"sdrxref_{kind}_gain").restype = C.c_float
        getattr(lib, f"sdrxref_{kind}_adjust").argtypes = [C.c_void_p, c_float_p, C.c_uint, c_float_p]
        getattr(lib, f"sdrxref_{kind}_adjust").restype = None
    lib.sdrxref_agc_set.argtypes = [C.c_void_p] + [C.c_float] * 4
    lib.sdrxref_agc_set.restype = None
    lib.sdrxref_lfagc_set.argtypes = [C.c_void_p] + [C.c_float] * 3 + [C.c_int]
    lib.sdrxref_lfagc_set.restype = None
    for kind in ("fir", "fir2", "fir3c"):
        getattr(lib, f"sdrxref_{kind}_create").argtypes = [c_float_p, C.c_int]
        getattr(lib, f"sdrxref_{kind}_create").restype = C.c_void_p
        getattr(lib, f"sdrxref_{kind}_destroy").argtypes = [C.c_void_p]
        getattr(lib, f"sdrxref_{kind}_destroy").restype = None
        getattr(lib, f"sdrxref_{kind}_set_gain").argtypes = [C.c_void_p, C.c_float]
        getattr(lib, f"sdrxref_{kind}_set_gain").restype = None
        getattr(lib, f"sdrxref_{kind}_filter").argtypes = [C.c_void_p, c_float_p, C.c_uint, c_float_p]
        getattr(lib, f"sdrxref_{kind}_filter").restype = None
    lib.sdrxref_rx_create.argtypes = [C.c_uint32, C.c_int, c_int_p, c_float_p, C.c_int, C.c_float, C.c_int, C.c_int]
    lib.sdrxref_rx_create.restype = C.c_void_p
    lib.sdrxref_rx_destroy.argtypes = [C.c_void_p]
    lib.sdrxref_rx_destroy.restype = None
    lib.sdrxref_rx_pos.argtypes = [C.c_void_p, C.c_int]
    lib.sdrxref_rx_pos.restype = C.c_int
    lib.sdrxref_rx_set_modulation.argtypes = [C.c_void_p, c_int_p]
    lib.sdrxref_rx_set_modulation.restype = None
    lib.sdrxref_ingest.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint]
    lib.sdrxref_ingest.restype = C.c_float
    lib.sdrxref_rx_channelize.argtypes = [C.c_void_p, c_float_p]
    lib.sdrxref_rx_channelize.restype = None
    lib.sdrxref_rx_demod.argtypes = [C.c_void_p, c_float_p, c_u8_p, c_float_p, c_float_p, c_float_p, c_i16_p]
    lib.sdrxref_rx_demod.restype = None
    lib.sdrxref_rx_gains.argtypes = [C.c_void_p, c_float_p, c_float_p]
    lib.sdrxref_rx_gains.restype = None
    lib.sdrxref_build_info.argtypes = []
    lib.sdrxref_build_info.restype = C.c_char_p


# ----------------------------------------------------------------------------- helpers

def _c64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.complex64)


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int32)


def _cf(a: np.ndarray):
    """complex64 array -> float* (interleaved re,im)."""
    return a.view(np.float32).ctypes.data_as(c_float_p)


Stages = Sequence[Tuple[int, Sequence[float]]]   # [(m, taps), ...] == std::vector<MSD::Stage>


def _pack_stages(stages: Stages):
    m = _i32([s[0] for s in stages])
    ntaps = _i32([len(s[1]) for s in stages])
    taps = _f32(np.concatenate([np.asarray(s[1], dtype=np.float32) for s in stages]))
    return m, ntaps, taps


def fft512(x) -> np.ndarray:
    """The oracle's FP32 radix-2 DIT 512-point forward DFT (oracle/fft512.c)."""
    lib = load_backend("oracle").lib
    xin = _c64(x)
    assert xin.shape == (512,)
    out = np.empty(512, dtype=np.complex64)
    lib.sdrx_oracle_fft512(_cf(xin), _cf(out))
    return out


# ----------------------------------------------------------------------------- tables (ref only)

RATES = (960000, 1200000, 1440000, 1600000, 1920000, 2400000, 2560000, 6000000, 10000000)
TABLE_NAMES = ("coeff_bp4am_channel", "passband_shape", "ramp_up", "ramp_down",
               "fs_00016_16bit_ch_amdemod_lpf1")


def ref_rate_config(rate: int, backend: str = "ref"):
    """(N, z, [(m, taps)...]) of /root/reference/src/sdrx.cpp:1584-1677, read out of the reference build."""
    lib = load_backend(backend).lib
    N, z, ns = C.c_int(), C.c_int(), C.c_int()
    m = np.zeros(8, dtype=np.int32)
    nt = np.zeros(8, dtype=np.int32)
    taps = np.zeros(4096, dtype=np.float32)
    rc = lib.sdrxref_rate_config(rate, C.byref(N), C.byref(z), C.byref(ns),
                                 m.ctypes.data_as(c_int_p), nt.ctypes.data_as(c_int_p), _fp(taps), taps.size)
    if rc != 0:
        raise ValueError(f"rate {rate} not supported by the reference (rc={rc})")
    stages, pos = [], 0
    for s in range(ns.value):
        stages.append((int(m[s]), taps[pos:pos + nt[s]].copy()))
        pos += int(nt[s])
    return N.value, z.value, stages


def ref_table(name: str, backend: str = "ref") -> np.ndarray:
    lib = load_backend(backend).lib
    n = lib.sdrxref_table(name.encode(), None, 0)
    if n < 0:
        raise KeyError(name)
    out = np.zeros(n, dtype=np.float32)
    lib.sdrxref_table(name.encode(), _fp(out), n)
    return out


def translator(N: int, z: int, ch_offset: int, backend: str = "oracle") -> np.ndarray:
    """sdrx.cpp:1682-1690.  Empty array when ch_offset == 0 (the reference leaves the vector empty)."""
    if ch_offset == 0:
        return np.zeros(0, dtype=np.complex64)
    be = load_backend(backend)
    out = np.zeros(N, dtype=np.complex64)
    (be.lib.sdrxref_translator if be.is_ref else be.lib.sdrxo_translator)(N, z, ch_offset, _cf(out))
    return out


def window(backend: str = "oracle") -> np.ndarray:
    be = load_backend(backend)
    out = np.zeros(512, dtype=np.float32)
    (be.lib.sdrxref_window if be.is_ref else be.lib.sdrxo_window)(_fp(out))
    return out


def audio_pos(ch: int, n: int, backend: str = "oracle") -> int:
    be = load_backend(backend)
    return int((be.lib.sdrxref_audio_pos if be.is_ref else be.lib.sdrxo_audio_pos)(ch, n))


def channel_to_offset(name: str, tuner_fq: int) -> int:
    return int(load_backend("oracle").lib.sdrxo_channel_to_offset(name.encode(), int(tuner_fq)))


def parse_fq(s: str, aeronautical: bool = True) -> int:
    return int(load_backend("oracle").lib.sdrxo_parse_fq(s.encode(), 1 if aeronautical else 0))


# ----------------------------------------------------------------------------- class mirrors

class OracleMSD:
    """msd.hpp `MSD`: MSD(translator, stages, use_ftfir), m(), decimate(in) -> out, copyable."""

    def __init__(self, translator_: Optional[np.ndarray], stages: Stages, use_ftfir: bool = False,
                 backend: str = "oracle", _handle=None):
        self.be = load_backend(backend)
        self._p = "sdrxref_msd_" if self.be.is_ref else "sdrxo_msd_"
        if _handle is not None:
            self.h = _handle
            return
        tr = _c64(translator_ if translator_ is not None else np.zeros(0, np.complex64))
        m, nt, taps = _pack_stages(stages)
        self.h = getattr(self.be.lib, self._p + "create")(
            _cf(tr) if tr.size else None, int(tr.size), len(stages),
            m.ctypes.data_as(c_int_p), nt.ctypes.data_as(c_int_p), _fp(taps), 1 if use_ftfir else 0)

    def clone(self) -> "OracleMSD":
        return OracleMSD(None, [], backend=self.be.name, _handle=getattr(self.be.lib, self._p + "clone")(self.h))

    @property
    def m(self) -> int:
        return int(getattr(self.be.lib, self._p + "m")(self.h))

    def decimate(self, x) -> np.ndarray:
        xin = _c64(x)
        out = np.empty(xin.size // 1 + 1, dtype=np.complex64)   # never more outputs than inputs
        n = getattr(self.be.lib, self._p + "decimate")(self.h, _cf(xin), xin.size, _cf(out))
        return out[:n].copy()

    def __del__(self):
        try:
            getattr(self.be.lib, self._p + "destroy")(self.h)
        except Exception:
            pass


class OracleAGC:
    """agc.hpp `AGC` (ctor defaults :30) + the setters sdrx.cpp:1702-1705 uses."""

    def __init__(self, attack=10.0, decay=0.01, reference=0.25, max_gain=200.0, alpha=0.01, backend="oracle"):
        self.be = load_backend(backend)
        if self.be.is_ref:
            self.h = self.be.lib.sdrxref_agc_create(attack, decay, reference, max_gain, alpha)
        else:
            self.s = _AgcS()
            self.be.lib.sdrxo_agc_init(C.byref(self.s), attack, decay, reference, max_gain, alpha)

    def set(self, attack, decay, reference, max_gain):
        if self.be.is_ref:
            self.be.lib.sdrxref_agc_set(self.h, attack, decay, reference, max_gain)
        else:
            self.s.reference, self.s.attack, self.s.decay = reference, attack, decay
            self.be.lib.sdrxo_agc_set_max_gain(C.byref(self.s), max_gain)

    @property
    def gain(self) -> float:
        return float(self.be.lib.sdrxref_agc_gain(self.h)) if self.be.is_ref else float(self.s.gain)

    def adjust(self, x) -> np.ndarray:
        xin = _c64(x)
        out = np.empty_like(xin)
        if self.be.is_ref:
            self.be.lib.sdrxref_agc_adjust(self.h, _cf(xin), xin.size, _cf(out))
        else:
            self.be.lib.sdrxo_agc_adjust(C.byref(self.s), _cf(xin), xin.size, _cf(out))
        return out

    def __del__(self):
        if getattr(self, "be", None) is not None and self.be.is_ref:
            self.be.lib.sdrxref_agc_destroy(self.h)


class OracleLfAGC:
    """agc.hpp `LfAGC` (:86)."""

    def __init__(self, attack=10.0, decay=0.01, reference=0.25, max_gain=200.0, alpha=0.005, backend="oracle"):
        self.be = load_backend(backend)
        if self.be.is_ref:
            self.h = self.be.lib.sdrxref_lfagc_create(attack, decay, reference, max_gain, alpha)
        else:
            self.s = _LfAgcS()
            self.be.lib.sdrxo_lfagc_init(C.byref(self.s), attack, decay, reference, max_gain, alpha)

    def set(self, attack, decay, reference, active):
        if self.be.is_ref:
            self.be.lib.sdrxref_lfagc_set(self.h, attack, decay, reference, 1 if active else 0)
        else:
            self.s.reference, self.s.attack, self.s.decay = reference, attack, decay
            self.s.active = 1 if active else 0

    @property
    def gain(self) -> float:
        return float(self.be.lib.sdrxref_lfagc_gain(self.h)) if self.be.is_ref else float(self.s.gain)

    def adjust(self, x) -> np.ndarray:
        xin = _f32(x)
        out = np.empty_like(xin)
        if self.be.is_ref:
            self.be.lib.sdrxref_lfagc_adjust(self.h, _fp(xin), xin.size, _fp(out))
        else:
            self.be.lib.sdrxo_lfagc_adjust(C.byref(self.s), _fp(xin), xin.size, _fp(out))
        return out

    def __del__(self):
        if getattr(self, "be", None) is not None and self.be.is_ref:
            self.be.lib.sdrxref_lfagc_destroy(self.h)


class OracleFIR:
    """fir.hpp: kind 1 = FIR, 2 = FIR2 (stereo interleaved), 3 = FIR3<iqsample_t>."""

    def __init__(self, kind: int, coeffs, backend="oracle"):
        self.be = load_backend(backend)
        self.kind = kind
        c = _f32(coeffs)
        if self.be.is_ref:
            self._n = {1: "fir", 2: "fir2", 3: "fir3c"}[kind]
            self.h = getattr(self.be.lib, f"sdrxref_{self._n}_create")(_fp(c), c.size)
        else:
            self.h = self.be.lib.sdrxo_fir_create(kind, _fp(c), c.size)

    def set_gain(self, db: float):
        if self.be.is_ref:
            getattr(self.be.lib, f"sdrxref_{self._n}_set_gain")(self.h, db)
        else:
            self.be.lib.sdrxo_fir_set_gain(self.h, db)

    def filter(self, x) -> np.ndarray:
        if self.kind == 3:
            xin = _c64(x)
            out = np.empty_like(xin)
            pin, pout, n = _cf(xin), _cf(out), xin.size
        else:
            xin = _f32(x)
            out = np.empty_like(xin)
            pin, pout, n = _fp(xin), _fp(out), xin.size
        if self.be.is_ref:
            getattr(self.be.lib, f"sdrxref_{self._n}_filter")(self.h, pin, n, pout)
        else:
            self.be.lib.sdrxo_fir_filter(self.h, pin, n, pout)
        return out

    def __del__(self):
        try:
            if self.be.is_ref:
                getattr(self.be.lib, f"sdrxref_{self._n}_destroy")(self.h)
            else:
                self.be.lib.sdrxo_fir_destroy(self.h)
        except Exception:
            pass


# ----------------------------------------------------------------------------- whole receiver

@dataclass
class BlockOut:
    """Everything one 32 ms block produces (sdrx.cpp data_cb + alsa_write_cb)."""
    pwr_dbfs: float
    ch_iq: np.ndarray      # [C,512] complex64   CRB chunk (sdrx.cpp:343-344)
    audio: np.ndarray      # [C,512] float32     per-channel sample that is mixed (0 where none)
    sql: np.ndarray        # [C] uint8           squelch state AFTER this block's decision
    snr: np.ndarray        # [C] float32
    mix_pre: np.ndarray    # [1024] float32      stereo mix before FIR2
    mix: np.ndarray        # [1024] float32      after FIR2
    pcm: np.ndarray        # [1024] int16


class _RxBase:
    C: int

    def ingest(self, raw: np.ndarray, fmt: int) -> float:
        raise NotImplementedError

    def channelize(self) -> np.ndarray:
        raise NotImplementedError

    def demod(self):
        raise NotImplementedError

    def process(self, raw: np.ndarray, fmt: int) -> BlockOut:
        pwr = self.ingest(raw, fmt)
        ch_iq = self.channelize()
        audio, sql, snr, mix_pre, mix, pcm = self.demod()
        return BlockOut(pwr, ch_iq, audio, sql, snr, mix_pre, mix, pcm)

    @staticmethod
    def _n_samples(raw: np.ndarray, fmt: int) -> int:
        if fmt == FMT_F32:
            if raw.dtype == np.complex64:
                return raw.size
            assert raw.dtype == np.float32
            return raw.size // 2
        assert raw.dtype == (np.uint8 if fmt == FMT_U8 else np.int16)
        return raw.size // 2


class RefRx(_RxBase):
    """The reference's own classes driven by the restated sdrx.cpp glue (oracle/ref_harness.cpp).

    ``threaded=True`` uses the reference's DS worker threads (sdrx.cpp ``-t``, ds.hpp)."""

    def __init__(self, rate: int, offsets: Sequence[int], sql_levels: Optional[Sequence[float]] = None,
                 use_lf_agc: bool = False, lf_gain_db: float = 0.0, use_ftfir: bool = False,
                 threaded: bool = False, backend: str = "ref", modulation: Optional[Sequence[int]] = None):
        self.be = load_backend(backend)
        assert self.be.is_ref
        self.C = len(offsets)
        off = _i32(offsets)
        sql = _f32(sql_levels) if sql_levels is not None else None
        self.h = self.be.lib.sdrxref_rx_create(rate, self.C, off.ctypes.data_as(c_int_p),
                                               _fp(sql) if sql is not None else None,
                                               1 if use_lf_agc else 0, lf_gain_db, 1 if use_ftfir else 0,
                                               1 if threaded else 0)
        if not self.h:
            raise ValueError(f"rate {rate} unsupported")
        if modulation is not None:          # 0 AM / 1 FM per channel (Channel::mod, sdrx.cpp:230)
            self.be.lib.sdrxref_rx_set_modulation(self.h, _i32(modulation).ctypes.data_as(c_int_p))

    def pos(self, ch: int) -> int:
        return int(self.be.lib.sdrxref_rx_pos(self.h, ch))

    def ingest(self, raw, fmt):
        raw = np.ascontiguousarray(raw)
        return float(self.be.lib.sdrxref_ingest(self.h, raw.ctypes.data_as(C.c_void_p), fmt, self._n_samples(raw, fmt)))

    def channelize(self):
        out = np.empty((self.C, CH_IQ_BUF_SIZE), dtype=np.complex64)
        self.be.lib.sdrxref_rx_channelize(self.h, _cf(out))
        return out

    def channelize_discard(self):
        self.be.lib.sdrxref_rx_channelize(self.h, None)

    def demod(self):
        audio = np.empty((self.C, CH_IQ_BUF_SIZE), dtype=np.float32)
        sql = np.empty(self.C, dtype=np.uint8)
        snr = np.empty(self.C, dtype=np.float32)
        mix_pre = np.empty(1024, dtype=np.float32)
        mix = np.empty(1024, dtype=np.float32)
        pcm = np.empty(1024, dtype=np.int16)
        self.be.lib.sdrxref_rx_demod(self.h, _fp(audio), sql.ctypes.data_as(c_u8_p), _fp(snr), _fp(mix_pre),
                                     _fp(mix), pcm.ctypes.data_as(c_i16_p))
        return audio, sql, snr, mix_pre, mix, pcm

    def gains(self):
        a = np.empty(self.C, dtype=np.float32)
        l = np.empty(self.C, dtype=np.float32)
        self.be.lib.sdrxref_rx_gains(self.h, _fp(a), _fp(l))
        return a, l

    def __del__(self):
        try:
            self.be.lib.sdrxref_rx_destroy(self.h)
        except Exception:
            pass


class OracleRx(_RxBase):
    """The plain-C restatement (oracle/sdrx_oracle.c) driven with explicit tables.

    ``tables`` carries ``N, z, stages, fir2_taps, passband_shape, ramp_up, ramp_down`` (the caller takes
    them from the reference build via :func:`ref_rate_config` / :func:`ref_table`, or from the extracted
    copy the product ships)."""

    def __init__(self, tables: dict, offsets: Sequence[int], sql_levels: Optional[Sequence[float]] = None,
                 pos: Optional[Sequence[int]] = None, use_lf_agc: bool = False, lf_gain_db: float = 0.0,
                 use_ftfir: bool = False, modulation: Optional[Sequence[int]] = None):
        self.be = load_backend("oracle")
        lib = self.be.lib
        self.C = len(offsets)
        N, z = int(tables["N"]), int(tables["z"])
        tr = np.zeros((self.C, N), dtype=np.complex64)
        has = np.zeros(self.C, dtype=np.uint8)
        for c, off in enumerate(offsets):
            if off != 0:
                lib.sdrxo_translator(N, z, int(off), _cf(tr[c]))
                has[c] = 1
        m, nt, taps = _pack_stages(tables["stages"])
        sql = _f32(sql_levels if sql_levels is not None else [9.0] * self.C)
        posv = _i32(pos if pos is not None else [lib.sdrxo_audio_pos(c, self.C) for c in range(self.C)])
        self._keep = [tr, has, m, nt, taps, sql, posv,
                      _f32(tables["fir2_taps"]), _f32(tables["passband_shape"]),
                      _f32(tables["ramp_up"]), _f32(tables["ramp_down"])]
        cfg = _RxConfig()
        cfg.n_channels, cfg.N = self.C, N
        cfg.translators, cfg.has_translator = _cf(tr), has.ctypes.data_as(c_u8_p)
        cfg.n_stages = len(tables["stages"])
        cfg.m, cfg.ntaps, cfg.taps = m.ctypes.data_as(c_int_p), nt.ctypes.data_as(c_int_p), _fp(taps)
        cfg.sql_level, cfg.pos = _fp(sql), posv.ctypes.data_as(c_int_p)
        cfg.use_lf_agc, cfg.use_ftfir, cfg.lf_gain_db = int(use_lf_agc), int(use_ftfir), lf_gain_db
        cfg.fir2_taps, cfg.fir2_ntaps = _fp(self._keep[7]), self._keep[7].size
        cfg.passband_shape, cfg.ramp_up, cfg.ramp_down = _fp(self._keep[8]), _fp(self._keep[9]), _fp(self._keep[10])
        self.h = lib.sdrxo_rx_create(C.byref(cfg))
        self.posv = posv
        if modulation is not None:          # 0 AM / 1 FM per channel (Channel::mod, sdrx.cpp:230)
            lib.sdrxo_rx_set_modulation(self.h, _i32(modulation).ctypes.data_as(c_int_p))

    def pos(self, ch: int) -> int:
        return int(self.posv[ch])

    def ingest(self, raw, fmt):
        raw = np.ascontiguousarray(raw)
        return float(self.be.lib.sdrxo_rx_ingest(self.h, raw.ctypes.data_as(C.c_void_p), fmt, self._n_samples(raw, fmt)))

    def channelize(self):
        out = np.empty((self.C, CH_IQ_BUF_SIZE), dtype=np.complex64)
        self.be.lib.sdrxo_rx_channelize(self.h, _cf(out))
        return out

    def channelize_discard(self):
        self.be.lib.sdrxo_rx_channelize(self.h, None)

    def demod(self):
        audio = np.empty((self.C, CH_IQ_BUF_SIZE), dtype=np.float32)
        sql = np.empty(self.C, dtype=np.uint8)
        snr = np.empty(self.C, dtype=np.float32)
        mix_pre = np.empty(1024, dtype=np.float32)
        mix = np.empty(1024, dtype=np.float32)
        pcm = np.empty(1024, dtype=np.int16)
        self.be.lib.sdrxo_rx_demod(self.h, _fp(audio), sql.ctypes.data_as(c_u8_p), _fp(snr), _fp(mix_pre),
                                   _fp(mix), pcm.ctypes.data_as(c_i16_p))
        return audio, sql, snr, mix_pre, mix, pcm

    def gains(self):
        a = np.empty(self.C, dtype=np.float32)
        l = np.empty(self.C, dtype=np.float32)
        self.be.lib.sdrxo_rx_gains(self.h, _fp(a), _fp(l))
        return a, l

    def __del__(self):
        try:
            self.be.lib.sdrxo_rx_destroy(self.h)
        except Exception:
            pass


def ref_tables(rate: int, backend: str = "ref") -> dict:
    """The ``tables`` dict :class:`OracleRx` takes, read out of the reference build."""
    N, z, stages = ref_rate_config(rate, backend)
    return dict(N=N, z=z, stages=stages, fir2_taps=ref_table("coeff_bp4am_channel", backend),
                passband_shape=ref_table("passband_shape", backend), ramp_up=ref_table("ramp_up", backend),
                ramp_down=ref_table("ramp_down", backend))


# ----------------------------------------------------------------------------- Airspy real -> IQ front end
# (SURVEY.md 8f rank 1).  "oracle": oracle/airspy_fe_oracle.c inside liboracle.so; "ref": the reference's own
# libairspy iqconverter_float.c + filters.h compiled into oracle/_ref/libairspy_ref.so.
_AIRSPY_REF = os.path.join(HERE, "_ref", "libairspy_ref.so")
_AIRSPY_LIB = {}


def airspy_ref_available() -> bool:
    return os.path.exists(_AIRSPY_REF)


def _airspy_ref():
    if "ref" not in _AIRSPY_LIB:
        if not os.path.exists(_AIRSPY_REF) and os.path.isdir(REF_SRC):
            build()
        lib = C.CDLL(_AIRSPY_REF)
        lib.sdrxref_hb_kernel_len.restype = C.c_int
        lib.sdrxref_hb_kernel.argtypes = [c_float_p]
        lib.sdrxref_iqconv_create.restype = C.c_void_p
        lib.sdrxref_iqconv_free.argtypes = [C.c_void_p]
        lib.sdrxref_iqconv_reset.argtypes = [C.c_void_p]
        lib.sdrxref_iqconv_process.argtypes = [C.c_void_p, c_float_p, C.c_int]
        _AIRSPY_LIB["ref"] = lib
    return _AIRSPY_LIB["ref"]


def _airspy_oracle():
    if "oracle" not in _AIRSPY_LIB:
        lib = load_backend("oracle").lib
        lib.sdrxo_airspy_fe_create.restype = C.c_void_p
        lib.sdrxo_airspy_fe_create.argtypes = [c_float_p, C.c_int]
        lib.sdrxo_airspy_fe_destroy.argtypes = [C.c_void_p]
        lib.sdrxo_airspy_fe_reset.argtypes = [C.c_void_p]
        lib.sdrxo_airspy_unpack12.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        lib.sdrxo_airspy_fe_process.argtypes = [C.c_void_p, C.c_void_p, C.c_int, c_float_p]
        lib.sdrxo_airspy_fe_avg.restype = C.c_float
        lib.sdrxo_airspy_fe_avg.argtypes = [C.c_void_p]
        _AIRSPY_LIB["oracle"] = lib
    return _AIRSPY_LIB["oracle"]


def ref_hb_kernel() -> np.ndarray:
    """HB_KERNEL_FLOAT (libairspy filters.h:28-30) out of the compiled reference header."""
    lib = _airspy_ref()
    out = np.empty(lib.sdrxref_hb_kernel_len(), np.float32)
    lib.sdrxref_hb_kernel(_fp(out))
    return out


def airspy_pack12(samples: np.ndarray) -> np.ndarray:
    """Inverse of airspy.c:323-338 unpack_samples: 8 twelve-bit samples -> three uint32 (test input maker)."""
    s = np.asarray(samples, np.uint32).reshape(-1, 8)
    a = (s[:, 0] << 20) | (s[:, 1] << 8) | (s[:, 2] >> 4)
    b = ((s[:, 2] & 0xf) << 28) | (s[:, 3] << 16) | (s[:, 4] << 4) | (s[:, 5] >> 8)
    c = ((s[:, 5] & 0xff) << 24) | (s[:, 6] << 12) | s[:, 7]
    return np.stack([a, b, c], axis=1).astype(np.uint32).ravel()


class OracleAirspyFE:
    """real uint16 ADC samples (0..4095) -> complex64 IQ at half the rate, state carried between calls."""

    def __init__(self, hb_kernel: np.ndarray, backend: str = "oracle"):
        self.backend = backend
        self.hb = _f32(hb_kernel)
        if backend == "ref":
            self.lib = _airspy_ref()
            assert np.array_equal(self.hb, ref_hb_kernel())
            self.h = self.lib.sdrxref_iqconv_create()
        else:
            self.lib = _airspy_oracle()
            self.h = self.lib.sdrxo_airspy_fe_create(_fp(self.hb), int(self.hb.size))

    def unpack12(self, packed: np.ndarray) -> np.ndarray:
        packed = np.ascontiguousarray(packed, np.uint32)
        n = packed.size // 3 * 8
        out = np.empty(n, np.uint16)
        _airspy_oracle().sdrxo_airspy_unpack12(packed.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), n)
        return out

    def process(self, raw_u16: np.ndarray) -> np.ndarray:
        raw = np.ascontiguousarray(raw_u16, np.uint16)
        assert raw.size % 4 == 0
        if self.backend == "ref":
            # airspy.c:311-321 convert_samples_float, then the converter in place
            buf = ((raw.astype(np.int32) - 2048).astype(np.float32) * np.float32(1.0 / 2048.0)).astype(np.float32)
            self.lib.sdrxref_iqconv_process(self.h, _fp(buf), int(buf.size))
            return buf[:raw.size].view(np.complex64)[:raw.size // 2].copy()
        out = np.empty(raw.size // 2, np.complex64)
        self.lib.sdrxo_airspy_fe_process(self.h, raw.ctypes.data_as(C.c_void_p), int(raw.size), _cf(out))
        return out

    def reset(self):
        (self.lib.sdrxref_iqconv_reset if self.backend == "ref" else self.lib.sdrxo_airspy_fe_reset)(self.h)

    def close(self):
        if self.h:
            (self.lib.sdrxref_iqconv_free if self.backend == "ref" else self.lib.sdrxo_airspy_fe_destroy)(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

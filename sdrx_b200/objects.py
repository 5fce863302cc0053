"""Python mirror of the reference's single-object class surface, backed by the CUDA library.

Same names, argument meaning and defaults as the reference's classes:

* :class:`MSD`   -- msd.hpp:54-75   ``MSD(translator, stages)``, ``m()``, ``decimate(in)``; copyable
* :class:`AGC`   -- agc.hpp:30-69   ``AGC(attack=10, decay=.01, reference=.25, max_gain=200, alpha=.01)``
* :class:`LfAGC` -- agc.hpp:86-131
* :class:`FIR` / :class:`FIR2` / :class:`FIR3` -- fir.hpp:78, :130, :28

The C++ drop-in headers with the identical surface are under ``include/sdrx_b200/``; these classes exist so the
parity tests read like tests of the reference's own classes.  All compute is on the GPU.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib
from ._lib import AgcState, SdrxError, check, c_float_p, c_int_p


def _fp(a: np.ndarray):
    return a.ctypes.data_as(c_float_p)


class MSD:
    def __init__(self, translator: Optional[np.ndarray], stages: Sequence[Tuple[int, np.ndarray]],
                 use_ftfir: bool = False, device: int = 0, _handle=None):
        self.lib = _lib.load()
        if _handle is not None:
            self.h = _handle
            return
        tr = None if translator is None or len(translator) == 0 else np.ascontiguousarray(translator, np.complex64)
        m = np.asarray([s[0] for s in stages], np.int32)
        nt = np.asarray([len(s[1]) for s in stages], np.int32)
        taps = np.concatenate([np.asarray(s[1], np.float32) for s in stages])
        h = C.c_void_p()
        check(self.lib.sdrx_msd_create(_fp(tr.view(np.float32)) if tr is not None else None,
                                       0 if tr is None else tr.size, len(stages), m.ctypes.data_as(c_int_p),
                                       nt.ctypes.data_as(c_int_p), _fp(taps), 1 if use_ftfir else 0, device, C.byref(h)))
        self.h = h

    def m(self) -> int:
        return int(self.lib.sdrx_msd_m(self.h))

    def decimate(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, np.complex64)
        out = np.empty(x.size // max(self.m(), 1) + 2, np.complex64)
        n = C.c_size_t()
        check(self.lib.sdrx_msd_decimate(self.h, _fp(x.view(np.float32)) if x.size else None, x.size,
                                         _fp(out.view(np.float32)), C.byref(n)))
        return out[:n.value].copy()

    def clone(self) -> "MSD":
        h = C.c_void_p()
        check(self.lib.sdrx_msd_clone(self.h, C.byref(h)))
        return MSD(None, [], _handle=h)

    def close(self):
        if getattr(self, "h", None):
            self.lib.sdrx_msd_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class AGC:
    _lf = False

    def __init__(self, attack=10.0, decay=0.01, reference=0.25, max_gain=200.0, alpha=None, device: int = 0):
        self.lib = _lib.load()
        self.device = device
        s = AgcState()
        s.attack, s.decay, s.reference = attack, decay, reference
        s.max_gain, s.min_gain = max_gain, np.float32(1.0) / np.float32(max_gain)
        s.alpha = (0.005 if self._lf else 0.01) if alpha is None else alpha
        s.gain = 1.0
        s.power = np.float32(reference) * np.float32(reference) if self._lf else reference   # agc.hpp:32, :89
        s.active = 1
        self.s = s

    def setAttack(self, v): self.s.attack = v        # noqa: E704
    def setDecay(self, v): self.s.decay = v          # noqa: E704
    def setReference(self, v): self.s.reference = v  # noqa: E704  (power_ keeps its constructed value)

    def setMaxGain(self, v):
        self.s.max_gain, self.s.min_gain = v, np.float32(1.0) / np.float32(v)

    def gain(self) -> float:
        return float(self.s.gain)

    def adjust(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, np.complex64)
        out = np.empty_like(x)
        check(self.lib.sdrx_agc_adjust(self.device, C.byref(self.s), _fp(x.view(np.float32)), x.size,
                                       _fp(out.view(np.float32))))
        return out

    def adjust_noalloc(self, x: np.ndarray) -> np.ndarray:
        """``adjust`` through sdrx_agc_adjust_noalloc: the fixed device array inside the library as scratch, no allocation;
        at most ``AGC_NOALLOC_MAX`` samples per call."""
        x = np.ascontiguousarray(x, np.complex64)
        out = np.empty_like(x)
        check(self.lib.sdrx_agc_adjust_noalloc(self.device, C.byref(self.s), _fp(x.view(np.float32)), x.size,
                                               _fp(out.view(np.float32))))
        return out


AGC_NOALLOC_MAX = 8192      # SDRX_AGC_NOALLOC_MAX (include/sdrx_b200.h)


class LfAGC(AGC):
    _lf = True

    def active(self) -> bool: return bool(self.s.active)   # noqa: E704
    def activate(self): self.s.active = 1                  # noqa: E704
    def deactivate(self): self.s.active = 0                # noqa: E704

    def adjust(self, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        out = np.empty_like(x)
        check(self.lib.sdrx_lfagc_adjust(self.device, C.byref(self.s), _fp(x), x.size, _fp(out)))
        return out


class _FirBase:
    kind = 1

    def __init__(self, coeffs: Sequence[float], device: int = 0):
        self.lib = _lib.load()
        c = np.ascontiguousarray(coeffs, np.float32)
        h = C.c_void_p()
        check(self.lib.sdrx_fir_create(self.kind, _fp(c), c.size, device, C.byref(h)))
        self.h = h
        self._gain = 0.0

    def setGain(self, db: float):
        check(self.lib.sdrx_fir_set_gain(self.h, db))
        self._gain = db

    def gain(self) -> float:
        return self._gain

    def filter(self, x: np.ndarray) -> np.ndarray:
        if self.kind == 3:
            x = np.ascontiguousarray(x, np.complex64)
            out = np.empty_like(x)
            check(self.lib.sdrx_fir_filter(self.h, _fp(x.view(np.float32)), x.size, _fp(out.view(np.float32))))
        else:
            x = np.ascontiguousarray(x, np.float32)
            out = np.empty_like(x)
            check(self.lib.sdrx_fir_filter(self.h, _fp(x), x.size, _fp(out)))
        return out

    def copy(self):
        """The reference's implicit copy (fir.hpp:69-74, :176-182): taps, gain and the delay line as it stands."""
        o = object.__new__(type(self))
        o.lib = self.lib
        h = C.c_void_p()
        check(self.lib.sdrx_fir_clone(self.h, C.byref(h)))
        o.h, o._gain = h, self._gain
        return o

    def close(self):
        if getattr(self, "h", None):
            self.lib.sdrx_fir_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class FIR(_FirBase):
    kind = 1


class FIR2(_FirBase):
    kind = 2


class FIR3(_FirBase):
    kind = 3


__all__ = ["MSD", "AGC", "LfAGC", "FIR", "FIR2", "FIR3", "SdrxError"]

"""Host-side mirror of the Airspy real -> IQ front end (include/sdrx_b200.h, csrc/airspy_fe.cu): what libairspy does on
the host for AIRSPY_SAMPLE_FLOAT32_IQ (airspy.c:393-396 -> unpack_samples, convert_samples_float,
iqconverter_float_process), as CUDA kernels, so the wire format is the ADC's 12-bit real samples."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib
from ._lib import FE_BESIDE_ENGINE, FE_EXACT_DC, FE_PACKED12, check
from .config import tables


def hb_kernel() -> np.ndarray:
    """HB_KERNEL_FLOAT (libairspy filters.h:28-30), shipped as data in sdrx_b200/data/tables.npz."""
    return np.ascontiguousarray(tables()["airspy_hb_kernel"], np.float32)


class AirspyFrontEnd:
    """``process(raw)``: n real ADC samples (uint16 0..4095, or uint32 words when ``packed``) -> n/2 complex64 IQ."""

    def __init__(self, max_samples: int = 640000, packed: bool = False, exact_dc: bool = False, device: int = 0,
                 kernel: Optional[np.ndarray] = None, beside_engine: bool = False):
        self.lib = _lib.load()
        self.packed = packed
        self.device = device
        k = hb_kernel() if kernel is None else np.ascontiguousarray(kernel, np.float32)
        h = C.c_void_p()
        flags = ((FE_PACKED12 if packed else 0) | (FE_EXACT_DC if exact_dc else 0)
                 | (FE_BESIDE_ENGINE if beside_engine else 0))
        check(self.lib.sdrx_airspy_fe_create(device, k.ctypes.data_as(_lib.c_float_p), int(k.size), int(max_samples), flags,
                                             C.byref(h)))
        self.h = h

    def _n(self, raw: np.ndarray) -> int:
        return raw.size // 3 * 8 if self.packed else raw.size

    def process(self, raw: np.ndarray, fetch: bool = True) -> Optional[np.ndarray]:
        raw = np.ascontiguousarray(raw, np.uint32 if self.packed else np.uint16)
        n = self._n(raw)
        out = np.empty(n // 2, np.complex64) if fetch else None
        check(self.lib.sdrx_airspy_fe_process(self.h, raw.ctypes.data_as(C.c_void_p), n,
                                              out.view(np.float32).ctypes.data_as(_lib.c_float_p) if fetch else None))
        self._keep = raw
        return out

    def process_device(self, d_ptr: int, n_samples: int) -> None:
        check(self.lib.sdrx_airspy_fe_process_device(self.h, C.c_void_p(d_ptr), n_samples))

    def feed(self, engine, raw: np.ndarray) -> None:
        """One buffer through the front end and on into ``engine`` (push_block_device), without a host sync."""
        ev = self.lib.sdrx_engine_ingest_event(engine.h)
        if ev:          # the engine has read the output of the call before last: that ring slot is free again
            check(self.lib.sdrx_airspy_fe_wait_event(self.h, C.c_void_p(ev)))
        self.process(raw, fetch=False)
        n = C.c_uint()
        d = self.lib.sdrx_airspy_fe_output(self.h, C.byref(n))
        check(self.lib.sdrx_engine_wait_event(engine.h, C.c_void_p(self.lib.sdrx_airspy_fe_done_event(self.h))))
        engine.push_device(d, _lib.FMT_F32, n.value)

    @property
    def output_ptr(self) -> int:
        return int(self.lib.sdrx_airspy_fe_output(self.h, None))

    @property
    def stream(self) -> int:
        return int(self.lib.sdrx_airspy_fe_stream(self.h))

    @property
    def launch_count(self) -> int:
        return int(self.lib.sdrx_airspy_fe_launch_count(self.h))

    def avg(self) -> float:
        v = C.c_float()
        check(self.lib.sdrx_airspy_fe_avg(self.h, C.byref(v)))
        return v.value

    def sync(self) -> None:
        check(self.lib.sdrx_airspy_fe_sync(self.h))

    def reset(self) -> None:
        check(self.lib.sdrx_airspy_fe_reset(self.h))

    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.sdrx_airspy_fe_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

"""Python host side of the B200 engine: a thin mirror of include/sdrx_b200.h.

:class:`Engine` is the receiver chain of one GPU -- what ``data_cb`` (sdrx.cpp:306) plus
``alsa_write_cb`` (sdrx.cpp:420) do per 32 ms block for every listened channel.  All compute happens
in the CUDA library; numpy arrays only carry bytes across the boundary.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import (FMT_F32, FMT_S16, FMT_U8, FLAG_FTFIR, FLAG_POLYPHASE, FLAG_TGEMM, FLAG_TGEMM_TC, FLAG_FORCE_GENERIC, FLAG_NO_DEMOD, FLAG_PARTIAL_MIX, FLAG_NO_PAIRING, CHUNK,
                   SdrxError,
                   check)
from .config import ChannelPlan, make_translator, tables


def fmt_of(raw: np.ndarray) -> int:
    if raw.dtype == np.uint8:
        return FMT_U8
    if raw.dtype == np.int16:
        return FMT_S16
    if raw.dtype in (np.float32, np.complex64):
        return FMT_F32
    raise TypeError(f"unsupported IQ dtype {raw.dtype}: use uint8, int16, float32 or complex64")


def n_samples_of(raw: np.ndarray) -> int:
    return raw.size if raw.dtype == np.complex64 else raw.size // 2


@dataclass
class BlockOut:
    pwr_dbfs: float
    ch_iq: Optional[np.ndarray]
    audio: np.ndarray
    sql: np.ndarray
    snr: np.ndarray
    mix_pre: np.ndarray
    mix: np.ndarray
    pcm: np.ndarray
    agc_gain: np.ndarray
    lf_gain: np.ndarray


class PinnedBuffer:
    """cudaMallocHost-backed numpy array (sdrx_host_alloc)."""

    def __init__(self, nbytes: int):
        self._lib = _lib.load()
        self.ptr = self._lib.sdrx_host_alloc(nbytes)
        if not self.ptr:
            raise SdrxError("sdrx_host_alloc failed: " + self._lib.sdrx_last_error().decode())
        self.nbytes = nbytes
        self.u8 = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(self.ptr))

    def view(self, dtype, count: Optional[int] = None) -> np.ndarray:
        a = self.u8.view(dtype)
        return a if count is None else a[:count]

    def free(self):
        if self.ptr:
            self._lib.sdrx_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Engine:
    def __init__(self, plan: ChannelPlan, device: int = 0, use_lf_agc: bool = False, lf_gain_db: float = 0.0,
                 block_len: Optional[int] = None, force_generic: bool = False, demod: bool = True,
                 translators: Optional[np.ndarray] = None, partial_mix: bool = False, no_pairing: bool = False,
                 use_ftfir: bool = False, n_streams: int = 1, polyphase: bool = False, table_gemm: bool = False,
                 tensor_cores: bool = False):
        """``n_streams`` > 1: that many independent IQ streams, all listening to ``plan``, in one engine (one launch
        sequence per block for all of them, BASELINE.json config 4): :meth:`push` then takes ``[n_streams][block]``
        and every per-channel result has ``n_streams * len(plan.offsets)`` entries, stream-major."""
        self.lib = _lib.load()
        self.plan = plan
        self.device = device
        self.NS = max(1, int(n_streams))
        self.Cs = len(plan.offsets)
        self.C = self.NS * self.Cs
        rc_ = plan.rate
        self.N = rc_.N
        self.block_len = int(block_len or rc_.block_len)
        t = tables()
        if translators is None:
            tr = np.zeros((self.Cs, self.N), dtype=np.complex64)
            for c, off in enumerate(plan.offsets):
                if off != 0:
                    tr[c] = make_translator(rc_.N, rc_.z, off)
        else:
            tr = np.ascontiguousarray(translators, dtype=np.complex64)
        has = np.asarray([1 if o != 0 else 0 for o in plan.offsets], dtype=np.uint8)
        m = np.asarray([s[0] for s in rc_.stages], dtype=np.int32)
        nt = np.asarray([len(s[1]) for s in rc_.stages], dtype=np.int32)
        taps = np.concatenate([s[1] for s in rc_.stages]).astype(np.float32)
        sql = np.asarray(plan.sql_levels, dtype=np.float32)
        pos = np.asarray(plan.pos, dtype=np.int32)
        mod = np.asarray(plan.modulation_codes(), dtype=np.uint8)
        fir2 = np.ascontiguousarray(t["coeff_bp4am_channel"], dtype=np.float32)
        shape = np.ascontiguousarray(t["passband_shape"], dtype=np.float32)
        rup = np.ascontiguousarray(t["ramp_up"], dtype=np.float32)
        rdn = np.ascontiguousarray(t["ramp_down"], dtype=np.float32)
        fp = lambda a: a.ctypes.data_as(_lib.c_float_p)   # noqa: E731
        cfg = _lib.EngineConfig()
        cfg.device = device
        cfg.n_channels = self.Cs
        cfg.n_streams = self.NS
        if translators is None and not has.any():
            # every channel sits on the tuner frequency: the reference builds no translator at all (empty vector,
            # sdrx.cpp:1684-1690) -- sdrx's canonical single-channel run
            cfg.translator_len = 0
            cfg.translators = None
            cfg.has_translator = None
        else:
            cfg.translator_len = self.N
            cfg.translators = tr.view(np.float32).ctypes.data_as(_lib.c_float_p)
            cfg.has_translator = has.ctypes.data_as(_lib.c_u8_p)
        cfg.n_stages = len(rc_.stages)
        cfg.m = m.ctypes.data_as(_lib.c_int_p)
        cfg.ntaps = nt.ctypes.data_as(_lib.c_int_p)
        cfg.taps = fp(taps)
        cfg.block_len = self.block_len
        cfg.sql_level = fp(sql)
        cfg.pos = pos.ctypes.data_as(_lib.c_int_p)
        cfg.modulation = mod.ctypes.data_as(_lib.c_u8_p)
        offs = np.asarray(plan.offsets, dtype=np.int32)
        cfg.ch_offset = offs.ctypes.data_as(_lib.c_int_p)
        cfg.use_lf_agc = 1 if use_lf_agc else 0
        cfg.lf_gain_db = lf_gain_db
        cfg.fir2_taps, cfg.fir2_ntaps = fp(fir2), fir2.size
        cfg.passband_shape, cfg.ramp_up, cfg.ramp_down = fp(shape), fp(rup), fp(rdn)
        cfg.flags = ((FLAG_FORCE_GENERIC if force_generic else 0) | (0 if demod else FLAG_NO_DEMOD)
                     | (FLAG_PARTIAL_MIX if partial_mix else 0) | (FLAG_NO_PAIRING if no_pairing else 0)
                     | (FLAG_FTFIR if use_ftfir else 0)      # --ftfir, sdrx.cpp:1199 / msd.hpp:104-131
                     | (FLAG_POLYPHASE if polyphase else 0)   # opt-in polyphase / DFT channelizer: not bit-compatible
                     | (FLAG_TGEMM if table_gemm else 0)      # opt-in fold + table GEMM instead of the time-domain channel bank
                     | (FLAG_TGEMM_TC if tensor_cores else 0))  # ... with the GEMM on tcgen05 (bf16 x 9 split operands)
        self.demod = demod
        h = C.c_void_p()
        check(self.lib.sdrx_engine_create(C.byref(cfg), C.byref(h)))
        self.h = h

    # ---- block interface
    def push(self, raw: np.ndarray, fmt: Optional[int] = None) -> None:
        """Queue one block (host array; pinned arrays from :class:`PinnedBuffer` avoid the staging copy)."""
        raw = np.ascontiguousarray(raw)
        check(self.lib.sdrx_engine_push_block(self.h, raw.ctypes.data_as(C.c_void_p),
                                              fmt_of(raw) if fmt is None else fmt, n_samples_of(raw) // self.NS))
        self._keep = raw      # a pinned source must stay alive until the copy has run

    def push_device(self, d_ptr: int, fmt: int, n_samples: int) -> None:
        check(self.lib.sdrx_engine_push_block_device(self.h, C.c_void_p(d_ptr), fmt, n_samples))

    def out_len(self) -> int:
        n = C.c_uint()
        check(self.lib.sdrx_engine_last_out_len(self.h, C.byref(n)))
        return n.value

    def channel_iq(self) -> np.ndarray:
        n = self.out_len()
        out = np.empty((self.C, n), dtype=np.complex64)
        check(self.lib.sdrx_engine_get_channel_iq(self.h, out.view(np.float32).ctypes.data_as(_lib.c_float_p)))
        return out

    def outputs(self, with_iq: bool = True) -> BlockOut:
        C_ = self.C
        audio = np.empty((C_, CHUNK), np.float32)
        sql = np.empty(C_, np.uint8)
        snr = np.empty(C_, np.float32)
        mix_pre = np.empty(self.NS * 2 * CHUNK, np.float32)
        mix = np.empty(self.NS * 2 * CHUNK, np.float32)
        pcm = np.empty(self.NS * 2 * CHUNK, np.int16)
        pwr = np.empty(self.NS, np.float32)
        ag = np.empty(C_, np.float32)
        lg = np.empty(C_, np.float32)
        o = _lib.BlockOutputs()
        fp = lambda a: a.ctypes.data_as(_lib.c_float_p)   # noqa: E731
        o.audio, o.sql_state, o.snr = fp(audio), sql.ctypes.data_as(_lib.c_u8_p), fp(snr)
        o.mix_pre, o.mix, o.pcm, o.pwr_dbfs = fp(mix_pre), fp(mix), pcm.ctypes.data_as(_lib.c_i16_p), fp(pwr)
        o.agc_gain, o.lf_gain = fp(ag), fp(lg)
        check(self.lib.sdrx_engine_get_outputs(self.h, C.byref(o)))
        return BlockOut(float(pwr[0]) if self.NS == 1 else pwr, self.channel_iq() if with_iq else None, audio, sql, snr, mix_pre, mix,
                        pcm, ag, lg)

    def make_output_buffers(self, pinned: bool = True, want: Sequence[str] = ("audio", "sql_state", "snr", "mix", "pcm", "pwr_dbfs")):
        """Reusable result buffers for :meth:`outputs_into` (pinned: D2H straight into them).  Returns
        ``(struct, arrays, nbytes)``; only the fields named in ``want`` are fetched."""
        C_, NS = self.C, self.NS
        spec = {"audio": (np.float32, C_ * CHUNK), "sql_state": (np.uint8, C_), "snr": (np.float32, C_),
                "mix_pre": (np.float32, NS * 2 * CHUNK), "mix": (np.float32, NS * 2 * CHUNK), "pcm": (np.int16, NS * 2 * CHUNK),
                "pwr_dbfs": (np.float32, NS), "agc_gain": (np.float32, C_), "lf_gain": (np.float32, C_),
                "audio_rows": (np.float32, C_ * CHUNK), "audio_chan": (np.int32, C_), "n_audio": (np.int32, NS)}
        o = _lib.BlockOutputs()
        arrays, keep, total = {}, [], 0
        for name in want:
            dt, n = spec[name]
            nbytes = n * np.dtype(dt).itemsize
            if pinned:
                pb = PinnedBuffer(nbytes)
                keep.append(pb)
                a = pb.view(dt, n)
            else:
                a = np.empty(n, dt)
            arrays[name] = a
            total += nbytes
            ptr_t = {np.float32: _lib.c_float_p, np.uint8: _lib.c_u8_p, np.int16: _lib.c_i16_p, np.int32: _lib.c_int_p}[dt]
            setattr(o, name, a.ctypes.data_as(ptr_t))
        arrays["_keep"] = keep
        return o, arrays, total

    @staticmethod
    def dense_audio(arrays, n_channels: int, n_streams: int = 1) -> np.ndarray:
        """Expand the sparse audio of :meth:`make_output_buffers` (``audio_rows`` / ``audio_chan`` / ``n_audio``) into
        the dense ``[C][512]`` array ``audio`` would have returned (closed channels are exact zeros).  ``n_channels``
        counts all streams; each stream's list and rows start ``n_channels / n_streams`` entries further."""
        out = np.zeros((n_channels, CHUNK), np.float32)
        per = n_channels // n_streams
        rows = arrays["audio_rows"].reshape(n_channels, CHUNK)
        for s in range(n_streams):
            n = int(arrays["n_audio"][s])
            out[s * per + arrays["audio_chan"][s * per:s * per + n]] = rows[s * per:s * per + n]
        return out

    def outputs_into(self, struct) -> None:
        check(self.lib.sdrx_engine_get_outputs(self.h, C.byref(struct)))

    def fetch_outputs(self, struct) -> None:
        """Queue the read-back of the last pushed block into (pinned) buffers and return at once."""
        check(self.lib.sdrx_engine_fetch_outputs(self.h, C.byref(struct)))

    def wait_outputs(self, back: int = 0) -> None:
        """Block until the fetch queued ``back`` pushes ago has landed in host memory."""
        check(self.lib.sdrx_engine_wait_outputs(self.h, back))

    def process(self, raw: np.ndarray, fmt: Optional[int] = None) -> BlockOut:
        self.push(raw, fmt)
        return self.outputs()

    def demod_block(self, ch_iq: np.ndarray, pwr_dbfs: float = 0.0) -> BlockOut:
        """Boundary B alone on caller-supplied channel IQ [C,512]."""
        x = np.ascontiguousarray(ch_iq, dtype=np.complex64)
        assert x.shape == (self.C, CHUNK)
        check(self.lib.sdrx_engine_demod_block(self.h, x.view(np.float32).ctypes.data_as(_lib.c_float_p), pwr_dbfs))
        return self.outputs(with_iq=False)

    def finish_mix(self, parts: np.ndarray):
        """Channel-sharded runs: add the per-GPU partial mixes ``parts[n][1024]`` in order, then FIR2 + s16 with
        this engine's carried FIR2 ring.  Returns ``(mix_pre, mix, pcm)``."""
        parts = np.ascontiguousarray(parts, np.float32).reshape(-1, 2 * CHUNK)
        mix_pre = np.empty(2 * CHUNK, np.float32)
        mix = np.empty(2 * CHUNK, np.float32)
        pcm = np.empty(2 * CHUNK, np.int16)
        check(self.lib.sdrx_engine_finish_mix(self.h, parts.ctypes.data_as(_lib.c_float_p), parts.shape[0],
                                              mix_pre.ctypes.data_as(_lib.c_float_p), mix.ctypes.data_as(_lib.c_float_p),
                                              pcm.ctypes.data_as(_lib.c_i16_p)))
        return mix_pre, mix, pcm

    def sync(self) -> None:
        check(self.lib.sdrx_engine_sync(self.h))

    def join(self) -> None:
        """Order :attr:`stream` behind the side stream's queued work (see sdrx_engine_join)."""
        check(self.lib.sdrx_engine_join(self.h))

    def reset(self) -> None:
        check(self.lib.sdrx_engine_reset(self.h))

    # ---- carried state
    def save_state(self) -> bytes:
        n = self.lib.sdrx_engine_state_size(self.h)
        buf = (C.c_uint8 * n)()
        check(self.lib.sdrx_engine_save_state(self.h, buf, n))
        return bytes(buf)

    def load_state(self, blob: bytes) -> None:
        buf = (C.c_uint8 * len(blob)).from_buffer_copy(blob)
        check(self.lib.sdrx_engine_load_state(self.h, buf, len(blob)))

    # ---- introspection
    @property
    def stream(self) -> int:
        return int(self.lib.sdrx_engine_stream(self.h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self.lib.sdrx_engine_launch_count(self.h))

    @property
    def uses_fast_path(self) -> bool:
        return bool(self.lib.sdrx_engine_uses_fast_path(self.h))

    @property
    def uses_table_gemm(self) -> bool:
        """Boundary A runs as fold + table GEMM (tgemm.cuh)."""
        return bool(self.lib.sdrx_engine_uses_table_gemm(self.h))

    @property
    def uses_tensor_cores(self) -> bool:
        """... and that GEMM runs on the tensor cores (tgemm_tc.cuh, tcgen05.mma)."""
        return bool(self.lib.sdrx_engine_uses_tensor_cores(self.h))

    @property
    def bank_pairs(self) -> int:
        """Conjugate (+-offset) channel pairs that share a lane slot of the specialised bank."""
        return int(self.lib.sdrx_engine_bank_pairs(self.h))

    def set_timing(self, on: bool) -> None:
        check(self.lib.sdrx_engine_set_timing(self.h, 1 if on else 0))

    def last_timing(self):
        a, b = C.c_float(), C.c_float()
        check(self.lib.sdrx_engine_last_timing(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def close(self) -> None:
        if getattr(self, "h", None):
            self.lib.sdrx_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

"""ctypes view of include/sdrx_b200.h.  Loading fails loudly: there is no CPU fallback."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SDRX_B200_LIB") or os.path.join(HERE, "libsdrx_b200.so")   # the override is for A/B runs of two builds

c_float_p = C.POINTER(C.c_float)
c_int_p = C.POINTER(C.c_int)
c_u8_p = C.POINTER(C.c_uint8)
c_i16_p = C.POINTER(C.c_int16)

FMT_U8, FMT_S16, FMT_F32 = 0, 1, 2
FLAG_FORCE_GENERIC, FLAG_NO_DEMOD, FLAG_PARTIAL_MIX, FLAG_NO_PAIRING, FLAG_FTFIR, FLAG_POLYPHASE, FLAG_TGEMM = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40
FLAG_TGEMM_TC = 0x80
CHUNK = 512


class EngineConfig(C.Structure):
    _fields_ = [
        ("device", C.c_int), ("n_channels", C.c_int), ("translator_len", C.c_int),
        ("translators", c_float_p), ("has_translator", c_u8_p),
        ("n_stages", C.c_int), ("m", c_int_p), ("ntaps", c_int_p), ("taps", c_float_p),
        ("block_len", C.c_uint), ("sql_level", c_float_p), ("pos", c_int_p),
        ("agc_attack", C.c_float), ("agc_decay", C.c_float), ("agc_reference", C.c_float),
        ("agc_max_gain", C.c_float), ("agc_alpha", C.c_float),
        ("lf_attack", C.c_float), ("lf_decay", C.c_float), ("lf_reference", C.c_float),
        ("lf_max_gain", C.c_float), ("lf_alpha", C.c_float),
        ("use_lf_agc", C.c_int), ("lf_gain_db", C.c_float),
        ("fir2_taps", c_float_p), ("fir2_ntaps", C.c_int),
        ("passband_shape", c_float_p), ("ramp_up", c_float_p), ("ramp_down", c_float_p),
        ("flags", C.c_uint),
        ("modulation", c_u8_p),
        ("ch_offset", c_int_p),
        ("n_streams", C.c_int),
    ]


class BlockOutputs(C.Structure):
    _fields_ = [
        ("audio", c_float_p), ("sql_state", c_u8_p), ("snr", c_float_p), ("mix_pre", c_float_p),
        ("mix", c_float_p), ("pcm", c_i16_p), ("pwr_dbfs", c_float_p), ("agc_gain", c_float_p),
        ("lf_gain", c_float_p), ("audio_rows", c_float_p), ("audio_chan", c_int_p), ("n_audio", c_int_p),
    ]


class AgcState(C.Structure):
    _fields_ = [(n, C.c_float) for n in
                ("attack", "decay", "reference", "min_gain", "max_gain", "alpha", "gain", "power")] + \
               [("active", C.c_int)]


# every symbol include/sdrx_b200.h declares: (name, restype, argtypes)
PROTOTYPES = [
    ("sdrx_last_error", C.c_char_p, []),
    ("sdrx_version", C.c_char_p, []),
    ("sdrx_device_count", C.c_int, []),
    ("sdrx_engine_create", C.c_int, [C.POINTER(EngineConfig), C.POINTER(C.c_void_p)]),
    ("sdrx_engine_destroy", None, [C.c_void_p]),
    ("sdrx_engine_push_block", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint]),
    ("sdrx_engine_push_block_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint]),
    ("sdrx_engine_last_out_len", C.c_int, [C.c_void_p, C.POINTER(C.c_uint)]),
    ("sdrx_engine_get_channel_iq", C.c_int, [C.c_void_p, c_float_p]),
    ("sdrx_engine_get_outputs", C.c_int, [C.c_void_p, C.POINTER(BlockOutputs)]),
    ("sdrx_engine_fetch_outputs", C.c_int, [C.c_void_p, C.POINTER(BlockOutputs)]),
    ("sdrx_engine_wait_outputs", C.c_int, [C.c_void_p, C.c_int]),
    ("sdrx_engine_demod_block", C.c_int, [C.c_void_p, c_float_p, C.c_float]),
    ("sdrx_engine_sync", C.c_int, [C.c_void_p]),
    ("sdrx_engine_join", C.c_int, [C.c_void_p]),
    ("sdrx_engine_finish_mix", C.c_int, [C.c_void_p, c_float_p, C.c_int, c_float_p, c_float_p, c_i16_p]),
    ("sdrx_engine_wait_stream", C.c_int, [C.c_void_p, C.c_void_p]),
    ("sdrx_engine_stream_wait_ingest", C.c_int, [C.c_void_p, C.c_void_p]),
    ("sdrx_engine_stream_wait_outputs", C.c_int, [C.c_void_p, C.c_void_p]),
    ("sdrx_engine_device_outputs", C.c_int, [C.c_void_p, C.c_int, C.POINTER(BlockOutputs)]),
    ("sdrx_engine_finish_mix_async", C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, c_float_p, c_float_p, c_i16_p]),
    ("sdrx_debug_timeline", C.c_int, [C.c_void_p, C.c_int]),
    ("sdrx_debug_timeline_read", C.c_int, [C.c_void_p, c_float_p, C.c_int]),
    ("sdrx_engine_state_size", C.c_size_t, [C.c_void_p]),
    ("sdrx_engine_save_state", C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    ("sdrx_engine_load_state", C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t]),
    ("sdrx_engine_reset", C.c_int, [C.c_void_p]),
    ("sdrx_engine_stream", C.c_void_p, [C.c_void_p]),
    ("sdrx_engine_launch_count", C.c_uint64, [C.c_void_p]),
    ("sdrx_engine_uses_fast_path", C.c_int, [C.c_void_p]),
    ("sdrx_engine_bank_pairs", C.c_int, [C.c_void_p]),
    ("sdrx_engine_uses_table_gemm", C.c_int, [C.c_void_p]),
    ("sdrx_engine_uses_tensor_cores", C.c_int, [C.c_void_p]),
    ("sdrx_engine_last_timing", C.c_int, [C.c_void_p, c_float_p, c_float_p]),
    ("sdrx_engine_set_timing", C.c_int, [C.c_void_p, C.c_int]),
    ("sdrx_debug_bank_plan_check", C.c_int, [C.c_int, C.c_int, C.c_int, c_int_p, c_int_p, c_int_p]),
    ("sdrx_debug_bank_plan_dump", C.c_int, [C.c_int, C.c_int, C.c_int, c_int_p, C.c_int]),
    ("sdrx_debug_bank_trace", C.c_int, [C.c_int]),
    ("sdrx_debug_bank_trace_read", C.c_int, [C.POINTER(C.c_uint64), C.c_int]),
    ("sdrx_debug_fe_trace", C.c_int, [C.c_int]),
    ("sdrx_debug_fe_trace_read", C.c_int, [C.POINTER(C.c_uint64)]),
    ("sdrx_host_alloc", C.c_void_p, [C.c_size_t]),
    ("sdrx_host_free", None, [C.c_void_p]),
    ("sdrx_device_alloc", C.c_void_p, [C.c_int, C.c_size_t]),
    ("sdrx_device_free", None, [C.c_int, C.c_void_p]),
    ("sdrx_device_upload", C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t]),
    ("sdrx_make_translator", None, [C.c_int, C.c_int, C.c_int, c_float_p]),
    ("sdrx_make_window", None, [c_float_p]),
    ("sdrx_msd_create", C.c_int, [c_float_p, C.c_int, C.c_int, c_int_p, c_int_p, c_float_p, C.c_int, C.c_int,
                                  C.POINTER(C.c_void_p)]),
    ("sdrx_msd_clone", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    ("sdrx_msd_destroy", None, [C.c_void_p]),
    ("sdrx_msd_m", C.c_size_t, [C.c_void_p]),
    ("sdrx_msd_decimate", C.c_int, [C.c_void_p, c_float_p, C.c_size_t, c_float_p, C.POINTER(C.c_size_t)]),
    ("sdrx_agc_adjust", C.c_int, [C.c_int, C.POINTER(AgcState), c_float_p, C.c_uint, c_float_p]),
    ("sdrx_lfagc_adjust", C.c_int, [C.c_int, C.POINTER(AgcState), c_float_p, C.c_uint, c_float_p]),
    ("sdrx_agc_adjust_noalloc", C.c_int, [C.c_int, C.POINTER(AgcState), c_float_p, C.c_uint, c_float_p]),
    ("sdrx_fir_create", C.c_int, [C.c_int, c_float_p, C.c_int, C.c_int, C.POINTER(C.c_void_p)]),
    ("sdrx_fir_clone", C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    ("sdrx_fir_destroy", None, [C.c_void_p]),
    ("sdrx_fir_set_gain", C.c_int, [C.c_void_p, C.c_float]),
    ("sdrx_fir_filter", C.c_int, [C.c_void_p, c_float_p, C.c_uint, c_float_p]),
    ("sdrx_airspy_fe_create", C.c_int, [C.c_int, c_float_p, C.c_int, C.c_uint, C.c_uint, C.POINTER(C.c_void_p)]),
    ("sdrx_airspy_fe_destroy", None, [C.c_void_p]),
    ("sdrx_airspy_fe_reset", C.c_int, [C.c_void_p]),
    ("sdrx_airspy_fe_process", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint, c_float_p]),
    ("sdrx_airspy_fe_process_device", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint]),
    ("sdrx_airspy_fe_output", C.c_void_p, [C.c_void_p, C.POINTER(C.c_uint)]),
    ("sdrx_airspy_fe_done_event", C.c_void_p, [C.c_void_p]),
    ("sdrx_airspy_fe_stream", C.c_void_p, [C.c_void_p]),
    ("sdrx_airspy_fe_sync", C.c_int, [C.c_void_p]),
    ("sdrx_airspy_fe_avg", C.c_int, [C.c_void_p, c_float_p]),
    ("sdrx_airspy_fe_launch_count", C.c_uint64, [C.c_void_p]),
    ("sdrx_engine_wait_event", C.c_int, [C.c_void_p, C.c_void_p]),
    ("sdrx_engine_ingest_event", C.c_void_p, [C.c_void_p]),
    ("sdrx_airspy_fe_wait_event", C.c_int, [C.c_void_p, C.c_void_p]),
]
FE_PACKED12, FE_EXACT_DC, FE_BESIDE_ENGINE = 0x1, 0x2, 0x4

_LIB = None


class SdrxError(RuntimeError):
    pass


def load() -> C.CDLL:
    """dlopen sdrx_b200/libsdrx_b200.so (built by ``python -m sdrx_b200.build``)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise SdrxError(f"{LIB_PATH} is missing: build the CUDA library with `python -m sdrx_b200.build` "
                        "(sdrx_b200 has no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, restype, argtypes in PROTOTYPES:
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _LIB = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise SdrxError(f"sdrx_b200 error {rc}: {load().sdrx_last_error().decode()}")

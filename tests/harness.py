"""Drivers shared by the tests: run a scenario through the CPU oracle / reference build and through
the CUDA engine, and compare block outputs."""
from __future__ import annotations

import hashlib
from typing import List, Optional

import numpy as np

from oracle import bindings as ob
from sdrx_b200 import config as cf
from tests.scenarios import Scenario

AUDIO_TOL = 1e-4        # BASELINE.json north_star: audio within 1e-4 relative RMS (FP32)


def product_tables(rate: int) -> dict:
    """The `tables` dict OracleRx takes, from the extracted copy the product ships."""
    rc = cf.rate_config(rate)
    t = cf.tables()
    return dict(N=rc.N, z=rc.z, stages=rc.stages, fir2_taps=t["coeff_bp4am_channel"],
                passband_shape=t["passband_shape"], ramp_up=t["ramp_up"], ramp_down=t["ramp_down"])


def make_cpu_rx(scn: Scenario, backend: str = "oracle", threaded: bool = False):
    if backend == "oracle":
        return ob.OracleRx(product_tables(scn.rate), scn.offsets, scn.sql_levels, None, scn.use_lf_agc, scn.lf_gain_db,
                           scn.use_ftfir, scn.mod_codes())
    return ob.RefRx(scn.rate, scn.offsets, scn.sql_levels, scn.use_lf_agc, scn.lf_gain_db, scn.use_ftfir, threaded, backend,
                    scn.mod_codes())


def run_cpu(scn: Scenario, backend: str = "oracle", blocks: Optional[List[np.ndarray]] = None) -> List[ob.BlockOut]:
    rx = make_cpu_rx(scn, backend)
    src = scn.source()
    out = []
    for b in range(scn.n_blocks):
        raw = blocks[b] if blocks is not None else src.block(b)
        out.append(rx.process(raw, scn.fmt))
    return out


def run_engine(scn: Scenario, force_generic: bool = False, blocks: Optional[List[np.ndarray]] = None, device: int = 0,
               no_pairing: bool = False, table_gemm: bool = False, tensor_cores: bool = False):
    from sdrx_b200 import Engine
    eng = Engine(scn.plan(), device=device, use_lf_agc=scn.use_lf_agc, lf_gain_db=scn.lf_gain_db,
                 force_generic=force_generic, no_pairing=no_pairing, use_ftfir=scn.use_ftfir, table_gemm=table_gemm,
                 tensor_cores=tensor_cores)
    src = scn.source()
    out = []
    for b in range(scn.n_blocks):
        raw = blocks[b] if blocks is not None else src.block(b)
        out.append(eng.process(raw, scn.fmt))
    fast = eng.uses_fast_path
    eng.close()
    return out, fast


def rel_rms(a: np.ndarray, b: np.ndarray) -> float:
    """||a-b|| / ||b|| over everything (b is the reference); 0 when both are all zero."""
    a = np.asarray(a).astype(np.complex128 if np.iscomplexobj(a) else np.float64).ravel()
    b = np.asarray(b).astype(np.complex128 if np.iscomplexobj(b) else np.float64).ravel()
    den = np.sqrt(np.sum(np.abs(b) ** 2))
    num = np.sqrt(np.sum(np.abs(a - b) ** 2))
    return 0.0 if den == 0.0 and num == 0.0 else float(num / max(den, 1e-30))


def squelch_margin(snr: np.ndarray, lvl: np.ndarray) -> float:
    """Smallest distance of any finite SNR from its two thresholds (lvl, lvl+3): near-ties are visible."""
    snr = np.asarray(snr, dtype=np.float64)
    lvl = np.asarray(lvl, dtype=np.float64)
    ok = np.isfinite(snr) & (lvl != 0.0)
    if not ok.any():
        return float("inf")
    return float(np.min(np.minimum(np.abs(snr - lvl), np.abs(snr - (lvl + 3.0)))[ok]))


def digest(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()

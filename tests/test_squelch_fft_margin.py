"""Row a20: the reference runs FFTW3f (sdrx.cpp:887-890, :538), which is absent here, and thresholds
its output (:563-573); the oracle -- and the CUDA detector, bit for bit -- run the plain radix-2 FFT of oracle/fft512.c.
This test pins the squelch DECISIONS against the choice of FFT: for every scenario the SNR of every channel-block is
recomputed from the same windowed chunk with other correct 512-point DFTs,

  * numpy's pocketfft in single precision (complex64 in, complex64 arithmetic) -- a different FP32 algorithm
    (mixed radix, other twiddle handling), i.e. what swapping the FFT library does;
  * a float64 DFT rounded once to float32 -- the best any FP32 FFT (FFTW3f included) can return,

and the largest |delta SNR| any of them causes must be (far) smaller than the smallest distance of any oracle SNR from
a threshold it is compared with (lvl, lvl + 3).  Then no correct FP32 FFT can flip a decision on these inputs.
"""
import numpy as np
import pytest

from oracle import bindings as ob
from sdrx_b200 import config as cf
from tests import scenarios as sc
from tests.harness import product_tables, run_cpu


def _snr_f32(X: np.ndarray, shape: np.ndarray) -> np.ndarray:
    """sdrx.cpp:542-566 on spectra X[C][512] (complex64): float32, every operation rounded, sums in the reference's
    order (vectorised over channels only)."""
    re, im = X.real.astype(np.float32), X.imag.astype(np.float32)
    norm = lambda r, i: r * r + i * i                                    # noqa: E731  float32 ops, rounded one by one
    sig = np.zeros(X.shape[0], np.float32)
    for i in range(3, 91):
        sig = sig + norm(re[:, i], im[:, i])
        sig = sig + norm(re[:, 512 - i], im[:, 512 - i])
    sig = sig / np.float32(176)
    hi = np.zeros(X.shape[0], np.float32)
    lo = np.zeros(X.shape[0], np.float32)
    for i in range(112, 157):
        hi = hi + norm(re[:, i] * shape[i], im[:, i] * shape[i])
        lo = lo + norm(re[:, 512 - i] * shape[512 - i], im[:, 512 - i] * shape[512 - i])
    noise = (hi / np.float32(45) + lo / np.float32(45)) / np.float32(2)
    with np.errstate(divide="ignore", invalid="ignore"):
        return (np.float32(10) * np.log10((sig / noise).astype(np.float32))).astype(np.float32)


def _ffts(x: np.ndarray):
    """x[C][512] complex64 -> {name: X[C][512] complex64} by three different correct DFTs."""
    oracle = np.stack([ob.fft512(r) for r in x])
    pocket32 = np.fft.fft(x.astype(np.complex64), axis=1)
    assert pocket32.dtype == np.complex64, "numpy >= 2 runs complex64 FFTs in single precision"
    exact = np.fft.fft(x.astype(np.complex128), axis=1).astype(np.complex64)
    return {"oracle radix-2": oracle, "pocketfft f32": pocket32, "f64 rounded once": exact}


def _long_run_sample():
    """The 12-channel sample of tests/test_gpu_parity.py::test_long_run..., whose smallest oracle margin (0.09 dB) is the
    tightest decision in the repo's scenarios."""
    sample = [-480, -479, -123, -37, -1, 0, 1, 37, 200, 333, 478, 479]
    key = {-479: lambda b: (b // 3) % 2 == 0, -37: lambda b: b % 7 < 4, 37: lambda b: b % 5 != 0, 0: lambda b: 10 <= b < 30,
           200: lambda b: b % 11 < 9, 479: lambda b: b >= 2}
    car = [sc.Carrier(o, 0.01 + 0.004 * i, 400.0 + 210.0 * i, 0.5, k) for i, (o, k) in enumerate(key.items())]
    car += [sc.Carrier(o, 0.03, 1000.0 + o, 0.4) for o in (-400, -250, 90, 410)]
    return sc.Scenario("long_run_sample", 10000000, sample, car, n_blocks=40, noise_sigma=0.0008, seed=4242)


SCENARIOS = [sc.config1(), sc.config2(n_blocks=10), sc.paired_channels(), sc.centre_channel()] + sc.config2_variants() \
    + sc.fm_scenarios()[:1] + [sc.config3(n_blocks=4, n_channels=96), _long_run_sample()]


@pytest.mark.parametrize("scn", SCENARIOS, ids=lambda s: s.name)
def test_no_correct_fp32_fft_can_flip_a_squelch_decision(scn):
    want = run_cpu(scn, "oracle")
    win = ob.window()
    shape = np.asarray(cf.tables()["passband_shape"], np.float32)
    lvl = np.asarray(scn.sql_levels if scn.sql_levels is not None else [9.0] * len(scn.offsets), np.float64)
    worst_delta, margin = 0.0, float("inf")
    for o in want:
        x = (o.ch_iq.astype(np.complex64) * win.astype(np.float32)[None, :]).astype(np.complex64)   # sdrx.cpp:533
        snr = {k: _snr_f32(X, shape) for k, X in _ffts(x).items()}
        ok = np.isfinite(o.snr)
        # the restated band sums reproduce the oracle's SNR from the oracle's FFT (log10 implementation aside)
        assert np.max(np.abs(snr["oracle radix-2"][ok] - o.snr[ok])) < 2e-5
        for k in ("pocketfft f32", "f64 rounded once"):
            worst_delta = max(worst_delta, float(np.max(np.abs(snr[k][ok].astype(np.float64) - o.snr[ok]))))
        live = ok & (lvl != 0.0)             # lvl == 0 opens whatever the SNR (sdrx.cpp:569): nothing to flip
        if live.any():
            d = np.minimum(np.abs(o.snr[live] - lvl[live]), np.abs(o.snr[live] - (lvl[live] + 3.0)))
            margin = min(margin, float(d.min()))
    print(f"{scn.name}: max |dSNR| over FFT implementations {worst_delta:.2e} dB, smallest threshold margin {margin:.3f} dB")
    assert worst_delta < 1e-3, "a correct FP32 FFT moved an SNR by more than a millidecibel"
    assert worst_delta * 10 < margin, f"FFT rounding ({worst_delta:.2e} dB) is not far below the tightest decision ({margin:.3f} dB)"

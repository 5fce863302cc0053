"""The synthetic workloads of BASELINE.json:configs as bench.py and the tests run them.

All of them are pure functions of their seeds, so any rank regenerates any block; nothing here computes the chain.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np

from .config import ChannelPlan
from .source import Carrier, SyntheticSource

RATE_10M = 10_000_000
EPOCH_BLOCKS = 4          # blocks a carrier's keying holds (bench ring: 8 distinct blocks = two keying epochs)


def _keyed(k: int, seed: int):
    """Carrier k is on or off for a whole epoch, pseudo-randomly (on ~70 % of the epochs); epoch = b // EPOCH_BLOCKS."""
    def on(b: int) -> bool:
        e = b // EPOCH_BLOCKS
        h = (1103515245 * (seed * 7919 + 31 * k + 104729 * e) + 12345) & 0x7FFFFFFF
        return (h >> 8) % 10 < 7
    return on


def config3(n_channels: int = 960, sql_level: Optional[float] = None, key_seed: int = 1) -> Tuple[ChannelPlan, List[Carrier], str]:
    """Config 3: Airspy 10 MS/s f32, offsets -C/2 .. C/2-1 (the whole 8 MHz band at C = 960), ~10 % of
    the channels carry a keyed AM transmitter (the carrier set of tests/scenarios.config3: amplitudes log-uniform,
    in-channel SNRs roughly 15..45 dB), keyed on/off pseudo-randomly per epoch.  ``sql_level = 0`` opens every squelch
    whatever the SNR (sdrx.cpp:569): the worst case of boundary B."""
    offs = list(range(-n_channels // 2, n_channels // 2))
    rng = np.random.Generator(np.random.PCG64(777))
    active = sorted(rng.choice(len(offs), size=max(1, len(offs) // 10), replace=False).tolist())
    carriers = []
    for k in active:
        amp = float(10 ** rng.uniform(-2.6, -1.3))
        _ = int(rng.integers(0, 2))                  # keeps the draw sequence of tests/scenarios.config3
        carriers.append(Carrier(offs[k], amp, float(rng.uniform(300, 3000)), 0.5, _keyed(k, key_seed)))
    lv = None if sql_level is None else [float(sql_level)] * len(offs)
    desc = (f"config3: Airspy 10 MS/s f32 IQ, full 8 MHz band, {n_channels} channels at 8.33 kHz spacing, "
            f"{len(carriers)} keyed AM carriers (~10 %), squelch+panorama")
    return ChannelPlan.from_offsets(RATE_10M, offs, lv), carriers, desc


def config2() -> Tuple[ChannelPlan, List[Carrier], str]:
    offs = [-360, -121, -37, -1, 2, 40, 250, 479]
    rng = np.random.Generator(np.random.PCG64(777))
    carriers = [Carrier(o, float(10 ** rng.uniform(-2.4, -1.3)), float(rng.uniform(300, 3000)), 0.5, _keyed(i, 1))
                for i, o in enumerate(offs)]
    return (ChannelPlan.from_offsets(RATE_10M, offs), carriers,
            "config2: Airspy 10 MS/s f32 IQ, 8 channels at 8.33 kHz designators, squelch+panorama")


def config4_stream(stream: int, n_channels: int = 128) -> Tuple[ChannelPlan, List[Carrier], str]:
    """Config 4: stream ``stream`` (seed = stream + 1) of 64 independent 10 MS/s streams, 128 channels
    on every 7th 8.33 kHz designator around the tuner centre."""
    offs = [7 * (i - n_channels // 2) for i in range(n_channels)]
    rng = np.random.Generator(np.random.PCG64(1000 + stream))
    active = sorted(rng.choice(len(offs), size=max(1, len(offs) // 10), replace=False).tolist())
    carriers = [Carrier(offs[k], float(10 ** rng.uniform(-2.4, -1.3)), float(rng.uniform(300, 3000)), 0.5, _keyed(k, stream + 1))
                for k in active]
    return (ChannelPlan.from_offsets(RATE_10M, offs), carriers,
            f"config4: 10 MS/s f32 stream {stream}, {n_channels} channels on every 7th 8.33 kHz designator")


def synth_blocks(carriers, n: int, seed: int, noise_sigma: float = 0.0008) -> List[np.ndarray]:
    src = SyntheticSource(RATE_10M, carriers, noise_sigma=noise_sigma, seed=seed, n_blocks=n)
    return [src.block(b) for b in range(n)]

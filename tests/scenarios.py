"""Seeded test scenarios shared by the oracle tests, the golden-vector generator and the GPU parity
tests (scaled so the CPU oracle finishes in seconds)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np

from sdrx_b200 import FMT_F32, FMT_S16, FMT_U8, Carrier, ChannelPlan, SyntheticSource


@dataclass
class Scenario:
    name: str
    rate: int
    offsets: List[int]
    carriers: List[Carrier]
    fmt: int = FMT_F32
    n_blocks: int = 8
    noise_sigma: float = 0.01
    seed: int = 12345
    sql_levels: Optional[List[float]] = None
    use_lf_agc: bool = False
    lf_gain_db: float = 0.0
    use_ftfir: bool = False         # --ftfir (sdrx.cpp:1199): msd.hpp:104-131
    modulation: Optional[List[str]] = None   # "AM"/"FM" per channel (sdrx.cpp:230); None -> AM

    def mod_codes(self) -> Optional[List[int]]:
        return None if self.modulation is None else [1 if m == "FM" else 0 for m in self.modulation]

    def plan(self) -> ChannelPlan:
        return ChannelPlan.from_offsets(self.rate, self.offsets, self.sql_levels, modulation=self.modulation)

    def source(self) -> SyntheticSource:
        return SyntheticSource(self.rate, self.carriers, self.noise_sigma, self.seed, self.fmt, self.n_blocks)


def _window(lo: int, hi: int) -> Callable[[int], bool]:
    return lambda b: lo <= b < hi


def config1() -> Scenario:
    """RTL-SDR 2.4 MS/s uint8, one channel on a 25 kHz designator (118.275 -> tuner 118.3 MHz, offset -3),
    AM 1 kHz tone keyed on for blocks 8..27: closed -> open (ramp up) -> closed (ramp down)."""
    return Scenario("config1_rtl_2400_u8", 2400000, [-3],
                    [Carrier(-3, 0.5, 1000.0, 0.5, _window(8, 28))], fmt=FMT_U8, n_blocks=40)


def config2(fmt: int = FMT_F32, n_blocks: int = 12) -> Scenario:
    """Airspy 10 MS/s, 8 channels spread over +-3 MHz with their own tones and keying schedules."""
    offs = [-360, -121, -37, -1, 2, 40, 250, 479]
    tones = [300.0, 700.0, 1000.0, 1300.0, 1700.0, 2100.0, 2600.0, 2400.0]
    amps = [0.05, 0.02, 0.1, 0.008, 0.2, 0.03, 0.012, 0.06]
    sched = [_window(1, 7), _window(0, 99), _window(3, 5), _window(2, 9), _window(4, 10), _window(0, 3),
             _window(6, 99), _window(2, 4)]
    return Scenario(f"config2_airspy_10000_{['u8', 's16', 'f32'][fmt]}", 10000000, offs,
                    [Carrier(o, a, t, 0.5, s) for o, a, t, s in zip(offs, amps, tones, sched)],
                    fmt=fmt, n_blocks=n_blocks, noise_sigma=0.002)


def config2_variants() -> List[Scenario]:
    a = config2()
    a.name += "_lfagc_vol"
    a.use_lf_agc, a.lf_gain_db, a.n_blocks = True, 6.0, 6
    b = config2()
    b.name += "_sql"
    b.sql_levels, b.n_blocks = [0.0, 20.0, 9.0, 3.0, 9.0, 12.0, 0.0, 30.0], 6
    return [a, b]


def ftfir_scenarios() -> List[Scenario]:
    """--ftfir (sdrx.cpp:1199, msd.hpp:104-131, :341-423): 10 MS/s f32 (incl. a centre channel, which keeps the
    no-translator path), 2.4 MS/s u8 and 2.56 MS/s (the rate whose N is not a multiple of m, msd.hpp:173)."""
    a = config2(n_blocks=5)
    a.offsets = a.offsets + [0]
    a.name += "_ftfir"
    b = config1()
    b.name += "_ftfir"
    b.n_blocks = 14
    c = [s for s in other_rates() if s.rate == 2560000][0]
    c.name += "_ftfir"
    for s in (a, b, c):
        s.use_ftfir = True
    return [a, b, c]


def fm_scenarios() -> List[Scenario]:
    """--modulation FM / NNN.NNN/SQL/FM (sdrx.cpp:1197, :1351-1358; Demod::demod :181-194): FM and AM channels mixed in
    one run, FM carriers keyed so that open -> ramp-down (envelope, :500) -> closed -> open again (prev_sample kept
    from the last open block) are all walked; one with the LfAGC active."""
    offs = [-300, -37, 2, 121, 400, 0]
    mods = ["FM", "AM", "FM", "FM", "AM", "FM"]
    car = [Carrier(-300, 0.08, 800.0, 0.0, _window(0, 3), fm_dev_hz=2500.0),
           Carrier(-37, 0.05, 1000.0, 0.5, _window(1, 99)),
           Carrier(2, 0.1, 1500.0, 0.0, _window(1, 99), fm_dev_hz=1200.0),
           Carrier(121, 0.03, 400.0, 0.2, lambda b: b in (0, 1, 4, 5), fm_dev_hz=3000.0),
           Carrier(400, 0.06, 2000.0, 0.5, _window(0, 4)),
           Carrier(0, 0.04, 600.0, 0.0, _window(2, 5), fm_dev_hz=800.0)]
    a = Scenario("fm_mixed_10000", 10000000, offs, car, n_blocks=7, noise_sigma=0.002, modulation=mods)
    b = Scenario("fm_mixed_10000_lfagc", 10000000, offs, car, n_blocks=7, noise_sigma=0.002, modulation=mods,
                 use_lf_agc=True, lf_gain_db=3.0)
    c = Scenario("fm_rtl_2400_u8", 2400000, [-3, 9], [Carrier(-3, 0.4, 1000.0, 0.0, _window(1, 6), fm_dev_hz=2000.0),
                                                       Carrier(9, 0.2, 700.0, 0.0, fm_dev_hz=1500.0)],
                 fmt=FMT_U8, n_blocks=8, modulation=["FM", "FM"])
    return [a, b, c]


def paired_channels(fmt: int = FMT_F32) -> Scenario:
    """+-offset channels whose translator tables are complex conjugates of each other share a lane slot of the
    specialised bank (bank_fast.cuh, NCH = 2); mixed here with a centre channel and unpaired ones."""
    offs = [-37, 37, -1, 1, 0, 250, -250, 5, -479, 479, 33]
    tones = [400.0, 900.0, 1300.0, 700.0, 1000.0, 2100.0, 1700.0, 500.0, 2600.0, 1500.0, 800.0]
    amps = [0.05, 0.02, 0.1, 0.03, 0.08, 0.04, 0.012, 0.06, 0.02, 0.09, 0.015]
    sched = [_window(1, 5), _window(0, 99), _window(2, 4), _window(0, 3), _window(1, 99), _window(3, 6), _window(0, 99),
             _window(2, 99), _window(0, 2), _window(1, 4), _window(4, 99)]
    return Scenario(f"paired_10000_{['u8', 's16', 'f32'][fmt]}", 10000000, offs,
                    [Carrier(o, a, t, 0.5, s) for o, a, t, s in zip(offs, amps, tones, sched)],
                    fmt=fmt, n_blocks=6, noise_sigma=0.002)


def centre_channel() -> Scenario:
    """A channel at the tuner centre takes the reference's no-translator path (msd.hpp:79-101)."""
    return Scenario("centre_10000", 10000000, [0, 5], [Carrier(0, 0.1, 800.0), Carrier(5, 0.05, 1200.0)],
                    n_blocks=3, noise_sigma=0.002)


def other_rates() -> List[Scenario]:
    """Every other sample rate the reference supports, two channels, three blocks (generic kernels)."""
    out = []
    for rate, span in ((960000, 40), (1200000, 50), (1440000, 60), (1600000, 70), (1920000, 90), (2560000, 120),
                       (6000000, 280)):
        offs = [-span, span // 3]
        out.append(Scenario(f"rate_{rate}", rate, offs,
                            [Carrier(offs[0], 0.2, 900.0), Carrier(offs[1], 0.1, 1500.0, 0.5, _window(1, 99))],
                            fmt=FMT_U8 if rate < 3000000 else FMT_S16, n_blocks=3, noise_sigma=0.005))
    return out


def config3(n_blocks: int = 3, n_channels: int = 960) -> Scenario:
    """Airspy 10 MS/s, the whole 8 MHz band: offsets -480..479, ~10 % of the channels keyed, amplitudes
    log-uniform so that in-channel SNRs span roughly 15..45 dB."""
    offs = list(range(-n_channels // 2, n_channels // 2))
    rng = np.random.Generator(np.random.PCG64(777))
    active = sorted(rng.choice(len(offs), size=max(1, len(offs) // 10), replace=False).tolist())
    carriers = []
    for i, k in enumerate(active):
        amp = float(10 ** rng.uniform(-2.6, -1.3))
        start = int(rng.integers(0, 2))
        carriers.append(Carrier(offs[k], amp, float(rng.uniform(300, 3000)), 0.5, _window(start, 99)))
    return Scenario(f"config3_band_{n_channels}", 10000000, offs, carriers, n_blocks=n_blocks, noise_sigma=0.0008)

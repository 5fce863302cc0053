#!/usr/bin/env python
"""Replay a capture through the B200 chain the way `sdrx` listens to a dongle: channels given like on the sdrx command
line (`NNN.NNN[/SQL[/AM|FM]]`, sdrx.cpp:1349-1382), the 16 kHz stereo s16 stream ALSA would get (sdrx.cpp:646-656,
:699) written to a WAV file, squelch openings / closings printed like the console line (sdrx.cpp:588-625).

  python tools/replay.py capture.wav 118.105 118.280/12 121.500 -o audio.wav
  python tools/replay.py rtl.bin --rate 2400000 --format u8 118.275
  python tools/replay.py adc_packed.bin --rate 10000000 --real packed12 122.450 122.455/9/FM

Host glue only: every sample is computed by libsdrx_b200.so on the GPU (no CPU path)."""
from __future__ import annotations

import argparse
import os
import struct
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

PCM_RATE = 16000            # sdrx.cpp:699


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("capture", help="raw or WAV capture (rtl_sdr / airspy_rx)")
    ap.add_argument("channels", nargs="+", help="NNN.NNN[/SQL[/AM|FM]] as on the sdrx command line")
    ap.add_argument("--rate", type=int, default=None, help="IQ sample rate (raw captures; WAV files carry it)")
    ap.add_argument("--format", choices=["u8", "s16", "f32"], default=None, help="raw IQ sample format")
    ap.add_argument("--real", choices=["u16", "s16", "packed12"], default=None,
                    help="the capture holds the Airspy's REAL ADC samples (airspy_rx -t 4 / 3 / 5 -p 1): run the front end")
    ap.add_argument("--tuner", type=float, default=None,
                    help="centre frequency of the capture in MHz (default: what sdrx would tune to for these channels)")
    ap.add_argument("--sql-level", type=float, default=9.0, help="default squelch level in dB (sdrx --sql-level)")
    ap.add_argument("--modulation", choices=["AM", "FM"], default="AM", help="default modulation (sdrx --modulation)")
    ap.add_argument("--volume", type=float, default=0.0, help="audio gain in dB (sdrx --volume)")
    ap.add_argument("--lf-agc", action="store_true", help="sdrx --lf-agc")
    ap.add_argument("--ftfir", action="store_true", help="sdrx --ftfir")
    ap.add_argument("-o", "--output", default=None, help="write the stereo s16 16 kHz audio to this WAV file")
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--quiet", action="store_true")
    return ap


def open_capture(args):
    from sdrx_b200 import FMT_F32, FMT_S16, FMT_U8, FileSource
    fmt = {None: None, "u8": FMT_U8, "s16": FMT_S16, "f32": FMT_F32}[args.format]
    kind = {None: None, "u16": "real_u16", "s16": "real_s16", "packed12": "real_packed12"}[args.real]
    return FileSource(args.capture, args.rate, fmt, kind=kind)


def pcm_wav_header(n_frames: int) -> bytes:
    data = n_frames * 4
    return (b"RIFF" + struct.pack("<I", 36 + data) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 2, PCM_RATE, PCM_RATE * 4, 4, 16)
            + b"data" + struct.pack("<I", data))


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    from sdrx_b200 import AirspyFrontEnd, ChannelPlan, Engine
    src = open_capture(args)
    plan = ChannelPlan.from_names(src.rate, args.channels, args.sql_level, args.modulation,
                                  None if args.tuner is None else int(round(args.tuner * 1e6)))
    if not args.quiet:
        print(f"{args.capture}: {src.n_blocks} blocks of {src.block_len} IQ samples at {src.rate} S/s ({src.kind}); "
              f"tuner {plan.tuner_fq / 1e6:.3f} MHz; channels "
              + ", ".join(f"{n}[{o:+d}]/{int(s)}/{m}" for n, o, s, m in zip(plan.names, plan.offsets, plan.sql_levels, plan.modulation)))
    eng = Engine(plan, device=args.device, use_lf_agc=args.lf_agc, lf_gain_db=args.volume, use_ftfir=args.ftfir)
    fe = AirspyFrontEnd(max_samples=2 * src.block_len, packed=src.packed, device=args.device, beside_engine=True) if src.kind != "iq" else None
    out = open(args.output, "wb") if args.output else None
    if out:
        out.write(pcm_wav_header(0))
    prev = np.zeros(len(plan.names), np.uint8)
    frames, openings = 0, 0
    t0 = time.perf_counter()
    for b, blk in enumerate(src):
        if fe is not None:
            fe.feed(eng, blk)
            o = eng.outputs(with_iq=False)
        else:
            eng.push(blk, src.fmt)
            o = eng.outputs(with_iq=False)
        if out:
            o.pcm.astype("<i2").tofile(out)
        frames += o.pcm.size // 2
        for c in np.flatnonzero(o.sql != prev):
            openings += int(o.sql[c])
            if not args.quiet:
                print(f"{b * 0.032:9.3f} s  {plan.names[c]}  {'open ' if o.sql[c] else 'close'}  snr {o.snr[c]:5.1f} dB  level {o.pwr_dbfs:6.1f} dBFS")
        prev = o.sql.copy()
    dt = time.perf_counter() - t0
    if out:
        out.seek(0)
        out.write(pcm_wav_header(frames))
        out.close()
    if not args.quiet:
        print(f"{src.n_blocks} blocks ({src.n_blocks * 0.032:.2f} s of band) in {dt:.2f} s; {openings} squelch openings; "
              f"{frames} audio frames" + (f" -> {args.output}" if args.output else ""))
    if fe is not None:
        fe.close()
    eng.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())

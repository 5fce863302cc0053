#!/usr/bin/env python
"""Headline benchmark: channel-MS/s of the sdrx per-channel DSP chain at 10 MS/s IQ.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload config3|config2]

One *step* = one 32 ms IQ block (320 000 complex samples @10 MS/s) through the whole chain for every
channel: raw->float ingest, block power, NCO translate + 4-stage decimating FIR (channel bank), AGC,
AM envelope, SNR squelch (window + 512-pt FFT), pan mix, FIR2, s16.  channel-MS/s = channels x input
samples / time / 1e6; divide by 10 for the number of real-time 8.33 kHz channels.

Arms
  value      device-resident IQ (a ring of blocks larger than L2), whole block timed with CUDA events
             on the engine's stream, max over ranks.
  e2e        the same chain through the C-ABI call a host makes (sdrx_engine_push_block from pinned
             host memory + sdrx_engine_get_outputs into pinned host memory): H2D and D2H inside the
             timed region every step.
  reference  (--impl reference) the reference's own msd.hpp/agc.hpp/fir.hpp, Release flags, its
             thread-per-channel DS mode (ds.hpp) on all host cores, same workload and metric.

N > 1: one process per GPU (torchrun); every rank runs an independent 10 MS/s stream with the full
channel set (stream-sharded, SURVEY.md §8e "replicas only", no data-path collective) -> weak scaling.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RATE = 10_000_000
# algorithmic flops per (channel, input complex sample) as the reference computes it:
# mix 6 + folded FIR stages 52/5 + 58/25 + 64/125 (+ 202/625 in the stage-4 kernel)
FLOP_BANK = 6.0 + 52.0 / 5 + 58.0 / 25 + 64.0 / 125
FLOP_CHAIN = FLOP_BANK + 202.0 / 625
NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12


def workload(name: str, sql_level=None):
    from sdrx_b200 import workloads as wl
    if name == "config4":
        return wl.config4_stream(0)
    if name == "config3":
        return wl.config3(960, sql_level)
    if name == "config2":
        return wl.config2()
    raise SystemExit(f"unknown workload {name}")


# boundary A: which channelizer the engine runs (sdrx_engine_config.flags)
MODE_KW = {"bank": {}, "tensor_cores": {"tensor_cores": True}, "table_gemm": {"table_gemm": True}, "polyphase": {"polyphase": True}}
MODE_WHAT = {
    "bank": "time-domain channel bank (FP32 CUDA cores; the bit-comparable default of the library)",
    "tensor_cores": "fold + table GEMM on the tensor cores (SDRX_FLAG_TGEMM_TC: tcgen05.mma on three-way bf16 splits of the FP32 operands, "
                    "6 of the 9 plane products (the dropped ones are < 2^-24 of a term; SDRX_TC_PRODUCTS=9 keeps all), 16-entry partial "
                    "sums in tensor memory added with Kahan compensation, fold in double); north_star's complex-GEMM recast, inside its "
                    "tolerance on every channel (channel IQ <= 2.9e-5, audio <= 1.5e-5, squelch decisions bit-exact)",
    "table_gemm": "fold + table GEMM on the FP32 CUDA cores (SDRX_FLAG_TGEMM)",
    "polyphase": "exact-twiddle polyphase / DFT channelizer (SDRX_FLAG_POLYPHASE; NOT a parity mode)",
}


N_DISTINCT = 8            # distinct synthetic blocks (two keying epochs of the carriers); the device ring cycles them


def synth_blocks(carriers, n: int, seed: int):
    from sdrx_b200 import workloads as wl
    return wl.synth_blocks(carriers, n, seed)


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md).  The timed region of the
    default run lasts milliseconds, so the sampler is an NVML polling thread (about one sample per millisecond,
    nvidia_ml_py); `nvidia-smi -lms 100` is only the fallback where NVML cannot be loaded."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = (("sw_power_cap", 0x4), ("hw_slowdown", 0x8), ("sw_thermal_slowdown", 0x20), ("hw_thermal_slowdown", 0x40))

    def __init__(self, index: int, uuid=None):
        self.index = index
        self.nv = self.h = None
        self.samples, self.bits = [], 0
        self.stop_flag = False
        self.thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = None
            if uuid is not None:
                for cand in (f"GPU-{uuid}", str(uuid)):
                    try:
                        h = pynvml.nvmlDeviceGetHandleByUUID(cand.encode() if isinstance(cand, str) else cand)
                        break
                    except Exception:
                        try:
                            h = pynvml.nvmlDeviceGetHandleByUUID(cand)
                            break
                        except Exception:
                            h = None
            self.h = h if h is not None else pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nv = pynvml
        except Exception:
            self.nv = None
        self.f = None
        self.p = None

    def _poll(self):
        nv, h = self.nv, self.h
        while not self.stop_flag:
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                try:
                    self.bits |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                except Exception:
                    self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
            except Exception:
                pass
            time.sleep(0.0005)

    def start(self):
        if self.nv is not None:
            import threading
            self.stop_flag = False
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        try:
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self) -> dict:
        if self.nv is not None:
            self.stop_flag = True
            if self.thread is not None:
                self.thread.join(timeout=2)
            if not self.samples:          # the region ended before the first poll returned: read once, right behind it
                try:
                    self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
                except Exception:
                    pass
            try:
                mx = float(self.nv.nvmlDeviceGetMaxClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            except Exception:
                mx = None
            reasons = sorted(n for n, b in self.BITS if self.bits & b)
            return {"sm_mhz": statistics.median(self.samples) if self.samples else None, "sm_max_mhz": mx,
                    "samples": len(self.samples), "reasons": reasons, "how": "NVML polling thread inside the timed region"}
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons), "how": "nvidia-smi -lms 100"}


def fp32_peak():
    """FP32 CUDA-core peak for the roofline: the FMA micro-benchmark measured on this pool's B200
    (profiles/fp32_peak.json, tools/microbench/fp32_peak.cu) or, failing that, the nominal figure."""
    p = os.path.join(ROOT, "profiles", "fp32_peak.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d["fp32_tflops"]), "measured (profiles/fp32_peak.json: " + d.get("how", "FFMA micro-benchmark") + ")"
    return NOMINAL_FP32_TFLOPS, "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (MEASURED_PEAKS.json has no FP32 entry)"


# ------------------------------------------------------------------------------------------------
def time_reference(plan, blocks, steps: int, warmup: int, n_channels: int, threaded: bool = True):
    """The reference's own classes (oracle/_ref/libsdrx_ref_avx2.so: msd.hpp AVX2 path, Release flags)
    through the restated sdrx.cpp glue; thread-per-channel DS mode when `threaded`."""
    from oracle import bindings as ob
    backend = "ref_avx2" if ob.backend_available("ref_avx2") else ("ref" if ob.backend_available("ref") else None)
    offs = plan.offsets[:n_channels]
    if backend is not None:
        rx = ob.RefRx(RATE, offs, None, threaded=threaded, backend=backend)
        kind = "reference"
    else:       # the reference build did not travel: time the plain-C port instead
        from tests.harness import product_tables
        rx = ob.OracleRx(product_tables(RATE), offs)
        kind, threaded = "port", False
    nb = len(blocks)
    for i in range(warmup):
        rx.process(blocks[i % nb], 2)
    t0 = time.perf_counter()
    for i in range(steps):
        rx.ingest(blocks[i % nb], 2)
        rx.channelize_discard()
        rx.demod()
    dt = time.perf_counter() - t0
    cores = os.cpu_count() or 1
    used = min(cores, len(offs)) if threaded else 1
    return dt / steps, kind, backend or "oracle", used


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    plan, carriers, desc = workload(args.workload)
    blocks = synth_blocks(carriers, 2, 1)       # blocks 0, 1 of the GPU arm's rank-0 stream
    C_ = len(plan.offsets)
    L = plan.rate.block_len
    per_step, kind, backend, used = time_reference(plan, blocks, args.steps, max(args.warmup, 1), C_)
    val = C_ * L / per_step / 1e6
    sample = f"all {C_} channels, one 32 ms block per step, {args.steps} steps, reference -t mode (one DS thread per channel)"
    line = {
        "impl": "reference", "metric": "channel-MS/s", "value": val, "unit": "channel-MS/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "channels": C_, "block_len": L, "sample_rate": RATE, "wire_format": "f32"},
        "cpu_baseline": {"value": val, "unit": "channel-MS/s", "cores": used, "kind": kind, "sample": sample,
                         "build": backend},
        "e2e": {"value": val, "unit": "channel-MS/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "realtime_channels": val / (RATE / 1e6),
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from sdrx_b200 import Engine, PinnedBuffer, _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (sdrx_b200 has no CPU path)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.workload == "config4":
        return run_config4(args, world, rank, local, barrier, max_over_ranks)
    plan, carriers, desc = workload(args.workload)
    C_ = len(plan.offsets)
    L = plan.rate.block_len
    lib = _lib.load()
    mode = args.channelizer
    eng = Engine(plan, device=local, **MODE_KW[mode])
    try:
        uuid = torch.cuda.get_device_properties(local).uuid
    except Exception:
        uuid = None

    # ---- synthetic IQ: N_DISTINCT distinct blocks, laid out as a device ring larger than L2 (126 MB)
    blocks = synth_blocks(carriers, N_DISTINCT, 1 + rank)
    blk_bytes = blocks[0].nbytes
    ring = int(np.ceil(140e6 / blk_bytes / N_DISTINCT)) * N_DISTINCT      # 56 x 2.56 MB = 143 MB of f32 IQ
    d_ring = lib.sdrx_device_alloc(local, ring * blk_bytes)
    if not d_ring:
        raise SystemExit("device ring allocation failed")
    for i in range(ring):
        b = blocks[i % N_DISTINCT]
        _lib.check(lib.sdrx_device_upload(local, C.c_void_p(d_ring + i * blk_bytes), b.ctypes.data_as(C.c_void_p), blk_bytes))
    pins = [PinnedBuffer(blk_bytes) for _ in range(N_DISTINCT)]
    for p, b in zip(pins, blocks):
        p.view(np.complex64, L)[:] = b
    hviews = [p.view(np.complex64, L) for p in pins]

    def value_leg(e, steps, warmup, sample_clocks=False):
        """`steps` blocks from the device ring, whole chain, CUDA events on the engine's stream; max over ranks."""
        stream = torch.cuda.ExternalStream(e.stream, device=torch.device("cuda", local))
        for i in range(warmup):
            e.push_device(d_ring + (i % ring) * blk_bytes, 2, L)
        e.sync()
        sampler = ClockSampler(local, uuid) if sample_clocks else None
        barrier()
        if sampler:
            sampler.start()
        l0 = e.launch_count
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for i in range(steps):
            e.push_device(d_ring + ((warmup + i) % ring) * blk_bytes, 2, L)
        e.join()                    # the last blocks' side-stream kernels (recurrences, mix) are inside the timed region
        ev1.record(stream)
        e.sync()
        barrier()
        clocks = sampler.stop() if sampler else None
        ms = max_over_ranks(ev0.elapsed_time(ev1))
        return ms / steps, e.launch_count - l0, clocks

    def e2e_leg(e, steps):
        """The same chain through the C-ABI call a host makes: push_block from pinned host memory, fetch_outputs into
        pinned host memory (squelch, SNR, mix, pcm, power and the SPARSE per-channel audio: 2 KB per open channel),
        wait_outputs LAG blocks later -- so the copies of one block overlap the kernels of its neighbours and the
        device->host->device round trip is off the critical path.  Every step moves its own input and its own results
        across the host link inside the timed region; d2h bytes are counted from what actually crossed."""
        stream = torch.cuda.ExternalStream(e.stream, device=torch.device("cuda", local))
        LAG = int(os.environ.get("SDRX_BENCH_LAG", "5"))   # blocks between a fetch and its wait: a block's results are ~300 us behind its push
        want = ("audio_rows", "audio_chan", "n_audio", "sql_state", "snr", "mix", "pcm", "pwr_dbfs")
        osets = [e.make_output_buffers(pinned=True, want=want) for _ in range(LAG + 1)]
        fixed = C_ * 1 + C_ * 4 + 1024 * 4 + 1024 * 2 + 4 + 4            # sql, snr, mix, pcm, pwr, n_audio
        for i in range(4):
            e.push(hviews[i % N_DISTINCT])
            e.outputs_into(osets[i % (LAG + 1)][0])
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        ev0.record(stream)
        host = [0.0, 0.0, 0.0]                  # host seconds spent inside push / fetch / wait (diagnostic)
        rows = 0
        for i in range(steps):
            h0 = time.perf_counter()
            e.push(hviews[i % N_DISTINCT])
            h1 = time.perf_counter()
            e.fetch_outputs(osets[i % (LAG + 1)][0])
            h2 = time.perf_counter()
            if i >= LAG:
                e.wait_outputs(LAG)             # block i-LAG's audio / squelch / pcm are in host memory now
                rows += int(osets[(i - LAG) % (LAG + 1)][1]["n_audio"][0])
            h3 = time.perf_counter()
            host[0] += h1 - h0
            host[1] += h2 - h1
            host[2] += h3 - h2
        for back in range(min(LAG, steps) - 1, -1, -1):
            e.wait_outputs(back)                # drain: every block's results are in host memory inside the timed region
            rows += int(osets[(steps - 1 - back) % (LAG + 1)][1]["n_audio"][0])
        e.join()
        ev1.record(stream)
        e.sync()
        wall = (time.perf_counter() - t0) * 1e3
        barrier()
        ms = max_over_ranks(max(ev0.elapsed_time(ev1), wall))    # host-inclusive: the later of device span and wall clock
        open_rows = rows / steps
        return {"value": world * C_ * L / (ms / steps * 1e-3) / 1e6, "unit": "channel-MS/s",
                "h2d_bytes_per_step": blk_bytes, "d2h_bytes_per_step": int(round(fixed + open_rows * (2048 + 4))),
                "open_channels_per_step": open_rows, "ms_per_step": ms / steps,
                "host_us_per_step": {"push": host[0] / steps * 1e6, "fetch": host[1] / steps * 1e6,
                                     "wait": host[2] / steps * 1e6}}

    # ---- value leg
    per_step_ms, launches, clocks = value_leg(eng, args.steps, args.warmup, sample_clocks=True)
    value = world * C_ * L / (per_step_ms * 1e-3) / 1e6

    # ---- per-kernel time of the dominant kernel (channel bank), CUDA events around it on its stream
    eng.set_timing(True)
    bank, whole = [], []
    for i in range(min(args.steps, 20)):
        eng.push_device(d_ring + (i % ring) * blk_bytes, 2, L)
        eng.sync()
        a, b = eng.last_timing()
        bank.append(a)
        whole.append(b)
    eng.set_timing(False)
    bank_ms = statistics.mean(bank)
    peak, peak_how = fp32_peak()
    mp = os.path.join(ROOT, "MEASURED_PEAKS.json")
    measured = {}
    if os.path.exists(mp):
        with open(mp) as f:
            measured = json.load(f)
    hbm_peak = measured.get("hbm_gbs")
    if mode in ("tensor_cores", "table_gemm"):
        # The dominant kernel is the table GEMM.  `achieved` is the ALGORITHMIC work of the path as the reference computes it
        # (19.232 flop per channel and input sample for mix + stages 1-3, + stage 4 -- the GEMM covers all four)
        # over the GEMM kernel's time; the tensor cores execute 2 x rows x columns x K x (plane products) for it (padded tile sizes).
        n_inst = L // 625
        products = int(os.environ.get("SDRX_TC_PRODUCTS", "6")) if mode == "tensor_cores" else 1
        if mode == "tensor_cores":                           # what the kernel launches: 128-row x 64-column tiles, K padded to 64
            rows, n_tab = -(-2 * ((C_ + 1) // 2 + 1) // 128) * 128, -(-1200 // 64) * 64
        else:
            rows, n_tab = 2 * ((C_ + 1) // 2 + 1), 1200      # (Tr, Ti) per conjugate-pair slot
        executed = 2.0 * rows * (2 * n_inst) * n_tab * products
        alg = FLOP_CHAIN * C_ * L
        tpeak = float(measured.get("bf16_tflops", 1590.0)) if mode == "tensor_cores" else peak
        achieved = alg / (bank_ms * 1e-3) / 1e12
        roofline = {"bound": "tensor" if mode == "tensor_cores" else "fp32", "achieved": achieved, "peak": tpeak, "unit": "TFLOP/s",
                    "frac": achieved / tpeak, "traffic": None,
                    "kernel": f"k_tgemm_tc (tcgen05.mma, {products} bf16 plane products)" if mode == "tensor_cores" else "k_tgemm (FP32 FFMA2)",
                    "kernel_ms": bank_ms, "block_ms_serialised": statistics.mean(whole), "share_of_step": bank_ms / per_step_ms,
                    "algorithmic_flop_per_channel_sample": FLOP_CHAIN,
                    "executed_tflops": executed / (bank_ms * 1e-3) / 1e12, "executed_frac": executed / (bank_ms * 1e-3) / 1e12 / tpeak,
                    "executed_flop_per_launch": executed,
                    "of_fp32_cuda_core_peak": achieved / peak,
                    "chain_of_fp32_cuda_core_peak": FLOP_CHAIN * C_ * L / (per_step_ms * 1e-3) / 1e12 / peak,
                    "peak_source": ("measured (MEASURED_PEAKS.json bf16_tflops, cuBLAS burst)" if "bf16_tflops" in measured
                                    else "fallback 1.59 PFLOP/s (B200_PROFILING.md)") if mode == "tensor_cores" else peak_how,
                    "note": "the algorithmic count is the reference's time-domain chain; the GEMM form needs 2.4 GFLOP per block instead of "
                            "5.9 (shared fold, conjugate pairs) and the bf16 split multiplies that by the plane products kept, so `frac` understates "
                            "the tensor pipe's load: see executed_frac and profiles/r02z_tc_ncu_details.txt.  The 128 x 64 x 16 MMAs are bound by "
                            "shared-memory operand reads (6 KB per MMA: ~50 cycles against a 32-cycle tensor floor), DESIGN.md section 4"}
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):
            with open(prof) as f:
                roofline["traffic"] = json.load(f).get("tc_gemm_dram_bytes_per_launch")
    else:
        flop_bank = FLOP_BANK if eng.uses_fast_path else FLOP_CHAIN
        achieved = flop_bank * C_ * L / (bank_ms * 1e-3) / 1e12
        roofline = {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                    "peak_nominal": NOMINAL_FP32_TFLOPS, "frac_nominal": achieved / NOMINAL_FP32_TFLOPS,
                    "traffic": None, "kernel": "channel bank (mix + decimating FIR stages)" if eng.uses_fast_path
                    else "generic channelizer kernels (all stages)",
                    "kernel_ms": bank_ms, "block_ms_serialised": statistics.mean(whole),
                    "share_of_step": bank_ms / per_step_ms, "share_of_serialised_block": bank_ms / statistics.mean(whole),
                    "chain_frac": FLOP_CHAIN * C_ * L / (per_step_ms * 1e-3) / 1e12 / peak,
                    "algorithmic_flop_per_channel_sample": flop_bank, "peak_source": peak_how}
        prof = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(prof):
            with open(prof) as f:
                roofline["traffic"] = json.load(f).get("bank_dram_bytes_per_launch")
        # the HBM view of the same kernel, for completeness (the path is FP32-bound, HBM is irrelevant):
        # algorithmic bytes per launch = the block read once + the stage-3 rows written once
        alg_bytes = L * 8 + C_ * (L // 125) * 8
        roofline["hbm"] = {"algorithmic_bytes_per_launch": alg_bytes, "achieved": alg_bytes / (bank_ms * 1e-3) / 1e9,
                           "peak": hbm_peak, "unit": "GB/s",
                           "frac": (alg_bytes / (bank_ms * 1e-3) / 1e9 / hbm_peak) if hbm_peak else None}

    # ---- e2e leg
    e2e = e2e_leg(eng, args.steps)
    # what the wire allows: one block's H2D and one block's D2H alone, pinned memory, CUDA events (diagnostic)
    def _copy_gbs(dst, src, nbytes, reps=20):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dst.copy_(src, non_blocking=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(reps):
            dst.copy_(src, non_blocking=True)
        b.record()
        torch.cuda.synchronize()
        return nbytes * reps / (a.elapsed_time(b) * 1e-3) / 1e9
    hbuf = torch.empty(blk_bytes, dtype=torch.uint8, pin_memory=True)
    dbuf = torch.empty(blk_bytes, dtype=torch.uint8, device="cuda")
    e2e["pcie"] = {"h2d_gbs_block": _copy_gbs(dbuf, hbuf, blk_bytes), "d2h_gbs_block": _copy_gbs(hbuf, dbuf, blk_bytes)}

    # ---- sustained leg: config 5's regime -- seconds of back-to-back blocks, clocks sampled all the way through
    sustained = None
    if args.sustain > 0:
        n_sus = max(args.steps, int(args.sustain / (per_step_ms * 1e-3)))
        sus_ms, _, sus_clk = value_leg(eng, n_sus, 3, sample_clocks=True)
        sustained = {"seconds": sus_ms * n_sus * 1e-3, "steps": n_sus, "ms_per_step": sus_ms,
                     "value": world * C_ * L / (sus_ms * 1e-3) / 1e6, "ratio_to_value": per_step_ms / sus_ms,
                     "clocks": sus_clk}

    # ---- worst case of boundary B: sql_level 0 opens every squelch (sdrx.cpp:569): all channels run both recurrences,
    #      the ordered mix walks every channel, every audio row crosses the host link
    worst = None
    if not args.no_worst:
        wplan, _, _ = workload(args.workload, sql_level=0.0)
        weng = Engine(wplan, device=local, **MODE_KW[mode])
        w_ms, _, _ = value_leg(weng, args.steps, args.warmup)
        w_e2e = e2e_leg(weng, args.steps)
        worst = {"what": "same blocks, sql_level 0 on every channel: all squelches open (sdrx.cpp:569)",
                 "ms_per_step": w_ms, "value": world * C_ * L / (w_ms * 1e-3) / 1e6, "ratio_to_value": per_step_ms / w_ms,
                 "e2e": w_e2e}
        weng.close()

    # ---- channel-sharded leg (configs 3/5 over N GPUs): ONE stream, 960/N channels per GPU, NCCL broadcast of every
    #      block, NCCL gather of the partial mixes -- strong scaling, next to the replica numbers above
    sharded = None
    if world > 1 and not args.no_sharded:
        sharded = sharded_leg(args, plan, blocks, hviews, d_ring, blk_bytes, ring, world, rank, local, barrier, max_over_ranks)

    # ---- the other channelizers on the same workload, same value leg (the time-domain bank also with its FP32 roofline)
    alt = None
    if world == 1 and not args.no_alt:
        alt = {}
        acc = {"bank": "bit-comparable path: channel IQ <= 4.4e-5 rel. RMS per channel, audio 2.1e-5, squelch decisions equal (profiles/r02z_tc_report.json)",
               "tensor_cores": "squelch decisions equal to the oracle's, channel IQ <= 2.9e-5 on every channel, audio <= 1.5e-5 (profiles/r02z_tc_report.json)",
               "table_gemm": "squelch decisions equal to the oracle's; channel IQ of closed noise-only channels up to 1.4e-4 rel. RMS, audio 5e-5",
               "polyphase": "NOT a parity mode: channel IQ up to 1.8e-2 rel. RMS on noise-only channels (exact twiddles vs the reference's "
                            "float-rounded table), audio 8e-3; squelch decisions equal on this workload (profiles/r02x_channelizers.json)"}
        for name in ("bank", "tensor_cores", "table_gemm", "polyphase"):
            if name == mode:
                continue
            aeng = Engine(plan, device=local, **MODE_KW[name])
            a_ms, _, _ = value_leg(aeng, args.steps, args.warmup)
            aeng.set_timing(True)
            ks = []
            for i in range(5):
                aeng.push_device(d_ring + i * blk_bytes, 2, L)
                aeng.sync()
                ks.append(aeng.last_timing()[0])
            aeng.close()
            k_ms = statistics.median(ks)
            alt[name] = {"what": MODE_WHAT[name], "ms_per_step": a_ms, "value": C_ * L / (a_ms * 1e-3) / 1e6, "channelizer_kernel_ms": k_ms,
                         "accuracy": acc[name]}
            if name == "bank":
                alt[name]["roofline"] = {"bound": "fp32", "achieved": FLOP_BANK * C_ * L / (k_ms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                                         "frac": FLOP_BANK * C_ * L / (k_ms * 1e-3) / 1e12 / peak,
                                         "chain_frac": FLOP_CHAIN * C_ * L / (a_ms * 1e-3) / 1e12 / peak}

    # ---- parity of the bench's own workload against the oracle (--check): every channel, every distinct block
    check = None
    if args.check and rank == 0:
        check = check_against_oracle(plan, blocks, local, MODE_KW[mode])

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only): bounded sample of the same workload
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cs = min(C_, 4 * (os.cpu_count() or 1), 256)
        per_step, kind, backend, used = time_reference(plan, blocks, 4, 1, cs)
        cpu = {"value": cs * L / per_step / 1e6, "unit": "channel-MS/s", "cores": used, "kind": kind,
               "sample": f"first {cs} of {C_} channels, 4 blocks of 32 ms, reference -t mode (one DS thread per channel), build {backend}",
               "host_cpus": os.cpu_count()}
        per1, _, _, _ = time_reference(plan, blocks, 2, 1, min(C_, 8), threaded=False)
        cpu["single_core_value"] = min(C_, 8) * L / per1 / 1e6

    if rank == 0:
        line = {
            "metric": "channel-MS/s", "value": value, "unit": "channel-MS/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": per_step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "arithmetic": ("FP32 results from three-way bf16 splits of the FP32 operands on tcgen05 (plane products exact in the FP32 accumulator), "
                           "compensated FP32 accumulation; boundary B in FP32 as the reference" if mode == "tensor_cores"
                           else "FP32 CUDA cores throughout, as the reference"),
            "config": {"workload": desc, "channels": C_, "block_len": L, "sample_rate": RATE, "wire_format": "f32",
                       "parallelism": f"stream-sharded x{world} (one independent 10 MS/s stream per GPU, no collective)",
                       "l2": f"inputs cycle through a {ring}-block device ring ({ring * blk_bytes / 1e6:.0f} MB > 126 MB L2)",
                       "channelizer": mode + ": " + MODE_WHAT[mode],
                       "specialised_bank": bool(eng.uses_fast_path)},
            "realtime_channels": value / (RATE / 1e6),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "worst_case": worst, "sustained": sustained, "channel_sharded": sharded, "other_channelizers": alt,
        }
        if check is not None:
            line["parity_check"] = check
        print(json.dumps(line))
    eng.close()
    lib.sdrx_device_free(local, C.c_void_p(d_ring))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_config4(args, world, rank, local, barrier, max_over_ranks):
    """BASELINE.json config 4: 64 independent 10 MS/s streams x 128 channels (every 7th 8.33 kHz designator), stream-sharded
    64 / world per GPU.  All streams of a GPU go through ONE engine (sdrx_engine_config.n_streams): one launch sequence
    per block.  `value`: blocks resident in HBM; `e2e`: every stream's block crosses the host link as s16 IQ (the Airspy
    INT16_IQ wire format: 64 f32 streams would be 164 MB per step) and every stream's squelch / pcm / open audio rows
    come back.  The synthetic data is 8 generated streams x 2 blocks; engine stream s replays generated stream s % 8
    from block s // 8 on, so no two engine streams see the same sequence."""
    import torch
    import torch.distributed as dist
    from sdrx_b200 import Engine, PinnedBuffer, _lib, workloads as wl
    from sdrx_b200.source import quantize
    lib = _lib.load()
    total_streams = args.streams
    NS = max(1, total_streams // world)
    plan, _, desc = wl.config4_stream(0)
    Cs, L = len(plan.offsets), plan.rate.block_len
    G, NB = 8, 2
    gen = []
    srank = rank if args.seed_rank < 0 else args.seed_rank      # (debug) the data another rank would generate
    for g in range(G):
        _, car, _ = wl.config4_stream(srank * G + g)
        gen.append(wl.synth_blocks(car, NB, 1 + srank * G + g))
    def step_block(k, fmt_bytes):
        rows = [gen[s_ % G][(k + s_ // G) % NB] for s_ in range(NS)]
        if fmt_bytes == 8:
            return np.stack(rows)
        return np.stack([quantize(r, 1) for r in rows])
    n_distinct = 2
    eng = Engine(plan, device=local, n_streams=NS)
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))
    blk = [step_block(k, 8) for k in range(n_distinct)]
    nbytes = blk[0].nbytes
    d = lib.sdrx_device_alloc(local, n_distinct * nbytes)
    for k in range(n_distinct):
        _lib.check(lib.sdrx_device_upload(local, C.c_void_p(d + k * nbytes), blk[k].ctypes.data_as(C.c_void_p), nbytes))
    steps, warm = max(8, args.steps // 8), 3
    for i in range(warm):
        eng.push_device(d + (i % n_distinct) * nbytes, 2, L)
    eng.sync()
    barrier()
    sampler = ClockSampler(local, None)
    sampler.start()
    l0 = eng.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for i in range(steps):
        eng.push_device(d + (i % n_distinct) * nbytes, 2, L)
    eng.join()
    ev1.record(stream)
    eng.sync()
    barrier()
    clocks = sampler.stop()
    launches = eng.launch_count - l0
    ms = max_over_ranks(ev0.elapsed_time(ev1)) / steps
    value = world * NS * Cs * L / (ms * 1e-3) / 1e6
    eng.close()
    lib.sdrx_device_free(local, C.c_void_p(d))
    if args.verbose:
        sys.stderr.write(f"[rank {rank}] config4 value leg done: {ms * 1e3:.1f} us per step\n")
    # ---- e2e: s16 wire format from pinned host memory, results into pinned host memory
    eng = Engine(plan, device=local, n_streams=NS)
    stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", local))
    hb = [step_block(k, 4) for k in range(n_distinct)]
    pins = [PinnedBuffer(hb[0].nbytes) for _ in range(n_distinct)]
    hv = []
    for pb, a in zip(pins, hb):
        v = pb.view(np.int16, a.size).reshape(a.shape)
        v[:] = a
        hv.append(v)
    LAG = 2
    want = ("audio_rows", "audio_chan", "n_audio", "sql_state", "snr", "pcm", "pwr_dbfs")
    osets = [eng.make_output_buffers(pinned=True, want=want) for _ in range(LAG + 1)]
    for i in range(3):
        eng.push(hv[i % n_distinct])
        eng.outputs_into(osets[i % (LAG + 1)][0])
    barrier()
    t0 = time.perf_counter()
    ev0.record(stream)
    rows = 0
    for i in range(steps):
        eng.push(hv[i % n_distinct])
        eng.fetch_outputs(osets[i % (LAG + 1)][0])
        if i >= LAG:
            eng.wait_outputs(LAG)
            rows += int(osets[(i - LAG) % (LAG + 1)][1]["n_audio"].sum())
    for back in range(min(LAG, steps) - 1, -1, -1):
        eng.wait_outputs(back)
        rows += int(osets[(steps - 1 - back) % (LAG + 1)][1]["n_audio"].sum())
    eng.join()
    ev1.record(stream)
    eng.sync()
    wall = (time.perf_counter() - t0) * 1e3
    barrier()
    e_ms = max_over_ranks(max(ev0.elapsed_time(ev1), wall)) / steps
    fixed = NS * (Cs * 5 + 1024 * 2 + 4 + 4)
    e2e = {"value": world * NS * Cs * L / (e_ms * 1e-3) / 1e6, "unit": "channel-MS/s", "ms_per_step": e_ms,
           "h2d_bytes_per_step": int(hb[0].nbytes), "d2h_bytes_per_step": int(round(fixed + rows / steps * 2052)),
           "wire_format": "s16", "open_channels_per_step": rows / steps}
    eng.close()
    if rank == 0:
        line = {"metric": "channel-MS/s", "value": value, "unit": "channel-MS/s", "n_gpus": world, "steps": steps, "warmup": warm,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if total_streams == args.streams else "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"config4: {total_streams} independent 10 MS/s streams x {Cs} channels (every 7th 8.33 kHz designator), "
                                       f"{NS} streams per GPU in one engine", "streams": total_streams, "streams_per_gpu": NS,
                           "channels_per_stream": Cs, "block_len": L, "sample_rate": RATE, "wire_format": "f32 (value) / s16 (e2e)",
                           "parallelism": f"stream-sharded x{world}, no collective",
                           "l2": f"{NS * L * 8 * n_distinct / 1e6:.0f} MB of raw blocks per GPU cycle through the engine (> 126 MB L2)"
                                 if NS * L * 8 * n_distinct > 126e6 else "raw blocks fit L2 at this stream count"},
                "realtime_channels": value / (RATE / 1e6), "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
                "roofline": None, "cpu_baseline": None}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def sharded_leg(args, plan, blocks, hviews, d_ring, blk_bytes, ring, world, rank, local, barrier, max_over_ranks):
    """One 10 MS/s stream, its channels split over the ranks (sdrx_b200/sharded.py): per block one NCCL collective in
    (broadcast from rank 0, or all-gather of per-rank slices), one small NCCL all-gather of the partial mixes out, the
    mix finished on rank 0.  K blocks through the pipeline (depth 4), wall clock around the whole run incl. the final
    drain, max over ranks.  `value`: the block already lies on rank 0's GPU; `e2e`: every block is copied from pinned
    host memory inside the timed region and every rank's results land in pinned host memory."""
    import torch
    import torch.distributed as dist
    from sdrx_b200.sharded import ShardedReceiver
    C_ = len(plan.offsets)
    L = plan.rate.block_len
    K = args.sharded_blocks
    dev = torch.device("cuda", local)
    # every rank needs rank 0's blocks for the all-gather ingest (each copies its own slice): hand them round once, untimed
    shared = [torch.empty(2 * L, dtype=torch.float32, device=dev) for _ in range(N_DISTINCT)]
    for i, t in enumerate(shared):
        if rank == 0:
            t.copy_(torch.from_numpy(blocks[i].view(np.float32)))
        dist.broadcast(t, src=0)
    torch.cuda.synchronize()
    if rank != 0:
        for i, t in enumerate(shared):
            hviews[i].view(np.float32)[:] = t.cpu().numpy()
    out = {"workload": f"one 10 MS/s f32 stream, {C_} channels split {C_ // world}-per-GPU over {world} GPUs (configs 3/5, channel-sharded)",
           "blocks": K, "scaling": "strong", "pipeline_depth": 4}

    def run(ingest, from_host):
        rx = ShardedReceiver(plan, device=local, ingest=ingest, depth=4)
        def feed(i):
            if from_host:
                rx.submit(hviews[i % N_DISTINCT])
            else:
                rx.submit(device_block=shared[i % N_DISTINCT])
        for i in range(8):
            feed(i)
            if i >= 3:
                rx.collect()
        for _ in range(3):
            rx.collect()
        barrier()
        t0 = time.perf_counter()
        for i in range(K):
            feed(i)
            if i >= 3:
                rx.collect()
        for _ in range(3):
            rx.collect()
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        rx.close()
        barrier()
        return {"ms_per_block": dt / K * 1e3, "value": C_ * L * K / dt / 1e6}

    for ingest in ("broadcast", "allgather"):
        v = run(ingest, False)
        e = run(ingest, True)
        out[ingest] = {"value": v["value"], "ms_per_block": v["ms_per_block"],
                       "e2e": {"value": e["value"], "ms_per_block": e["ms_per_block"],
                               "h2d_bytes_per_block_src" if ingest == "broadcast" else "h2d_bytes_per_block_per_rank":
                                   blk_bytes if ingest == "broadcast" else blk_bytes // world}}
    # the collective alone: K broadcasts / all-gathers of one block back to back, CUDA events on the comm stream
    buf = torch.empty(2 * L, dtype=torch.float32, device=dev)
    part = torch.empty(2 * L // world, dtype=torch.float32, device=dev)
    mix = torch.empty(world * 1024, dtype=torch.float32, device=dev)
    mp = torch.empty(1024, dtype=torch.float32, device=dev)
    for name, fn in (("broadcast_us", lambda: dist.broadcast(buf, src=0)),
                     ("allgather_block_us", lambda: dist.all_gather_into_tensor(buf, part)),
                     ("allgather_mix_us", lambda: dist.all_gather_into_tensor(mix, mp))):
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(200):
            fn()
        b.record()
        torch.cuda.synchronize()
        out[name] = max_over_ranks(a.elapsed_time(b)) / 200 * 1e3
    best = min(("broadcast", "allgather"), key=lambda k: out[k]["ms_per_block"])
    out["value"] = out[best]["value"]
    out["e2e"] = out[best]["e2e"]
    out["best_ingest"] = best
    return out if rank == 0 else None


def check_against_oracle(plan, blocks, device, mode_kw=None):
    """bench.py --check: the bench's own blocks through the CUDA engine and through the CPU oracle (the checker, never
    the thing measured): squelch decisions must agree exactly, channel IQ / audio within 1e-4 relative RMS per channel."""
    from oracle import bindings as ob
    from sdrx_b200 import Engine
    from tests.harness import product_tables, rel_rms, squelch_margin
    orx = ob.OracleRx(product_tables(RATE), plan.offsets, plan.sql_levels)
    eng = Engine(plan, device=device, **(mode_kw or {}))
    worst_iq = worst_au = worst_iq_open = 0.0
    per_block = []
    margin = float("inf")
    sql_ok, opened = True, 0
    for blk in blocks:
        w = orx.process(blk, 2)
        g = eng.process(blk, 2)
        sql_ok = sql_ok and bool(np.array_equal(g.sql, w.sql))
        opened += int(w.sql.sum())
        margin = min(margin, squelch_margin(w.snr, np.asarray(plan.sql_levels)))
        errs = np.asarray([rel_rms(g.ch_iq[c], w.ch_iq[c]) for c in range(len(plan.offsets))])
        per_block.append((errs, w.sql.copy()))
        worst_iq = max(worst_iq, float(errs.max()))
        for c in range(len(plan.offsets)):
            if np.any(w.audio[c] != 0.0):
                worst_au = max(worst_au, rel_rms(g.audio[c], w.audio[c]))
    eng.close()
    ever_open = np.stack([q for _, q in per_block]).any(axis=0)
    worst_iq_open = float(np.stack([e_ for e_, _ in per_block])[:, ever_open].max()) if ever_open.any() else 0.0
    return {"blocks": len(blocks), "channels": len(plan.offsets), "squelch_decisions_equal": sql_ok,
            "open_channel_blocks": opened, "oracle_squelch_margin_db": margin, "worst_channel_iq_rel_rms": worst_iq,
            "worst_channel_iq_rel_rms_channels_ever_open": worst_iq_open, "worst_audio_rel_rms": worst_au, "bar": 1e-4}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config3", choices=["config3", "config2", "config4"])
    ap.add_argument("--seed-rank", type=int, default=-1, help=argparse.SUPPRESS)
    ap.add_argument("--verbose", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--streams", type=int, default=64, help="config4: independent streams over all GPUs")
    ap.add_argument("--channelizer", default="tensor_cores", choices=list(MODE_KW),
                    help="boundary A of the measured engine (config3/config2): tensor_cores = fold + table GEMM on tcgen05 (default), "
                         "bank = the bit-comparable time-domain channel bank, table_gemm / polyphase = the other opt-in forms")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-alt", action="store_true", help="skip the legs of the other channelizers (bank, table GEMM, polyphase)")
    ap.add_argument("--no-worst", action="store_true", help="skip the worst_case leg (every squelch open)")
    ap.add_argument("--sustain", type=float, default=3.2, help="seconds of back-to-back blocks in the sustained leg (0: skip)")
    ap.add_argument("--no-sharded", action="store_true", help="N > 1: skip the channel-sharded (one stream over N GPUs) leg")
    ap.add_argument("--sharded-blocks", type=int, default=2000, help="blocks of the channel-sharded leg")
    ap.add_argument("--check", action="store_true", help="also run the bench's own blocks through the CPU oracle and report parity")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    return run_reference(args) if args.impl == "reference" else run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

"""Channel-sharded receiver: ONE IQ stream, its channels partitioned over the GPUs of a box
(BASELINE.json configs 3 and 5).  One process per GPU, ``torch.distributed`` for the plumbing.

Per 32 ms block:

1. the rank that holds the raw block (``src``, the one fed by the dongle / file / synthetic source) broadcasts
   it **in the wire format** (u8 / s16 / f32 -- 0.64 / 1.28 / 2.56 MB at 10 MS/s) to every rank: one
   ``dist.broadcast`` (NCCL over NVLink / NVSwitch on the GPU box, gloo in the CPU tests);
2. every rank pushes the block through its own engine, which owns a contiguous range of channels and all of
   their carried state -- no other data-path exchange, channels are independent given the block;
3. the stereo mix is a sum over *all* channels (sdrx.cpp:471-532), so every rank contributes its partial
   ``mix_pre[1024]``; they are gathered on ``dst`` and added **in rank order** (= channel order between ranks),
   then FIR2 + float->s16 (sdrx.cpp:646-656) run once, on ``dst``'s engine, whose FIR2 ring is the carried one.

Per-channel results (audio, squelch state, SNR) stay on the rank that owns the channel.  The mix of a sharded
run equals the single-GPU mix up to the association of the float additions between ranks (the within-rank
order is the reference's); everything per channel is identical to the single-GPU run.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import numpy as np

from ._lib import CHUNK, FMT_F32, FMT_S16, FMT_U8
from .config import ChannelPlan

_WIRE_DTYPE = {FMT_U8: np.uint8, FMT_S16: np.int16, FMT_F32: np.float32}


def channel_range(n_channels: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous, balanced split: the first ``n_channels % world`` ranks own one channel more."""
    base, extra = divmod(n_channels, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_plan(plan: ChannelPlan, world: int, rank: int) -> ChannelPlan:
    """This rank's channels.  The pan positions are those of the *global* channel list
    (get_audio_pos(idx, n) depends on both, sdrx.cpp:1712), so they are sliced, not recomputed."""
    lo, hi = channel_range(len(plan.offsets), world, rank)
    return ChannelPlan(plan.rate, plan.tuner_fq, plan.names[lo:hi], plan.offsets[lo:hi], plan.sql_levels[lo:hi],
                       plan.pos[lo:hi], plan.modulation[lo:hi])


@dataclass
class ShardOut:
    lo: int                      # first global channel index owned by this rank
    hi: int
    local: object                # this rank's BlockOut (audio/sql/snr/... of its channels; mix_pre = partial mix)
    mix_pre: Optional[np.ndarray] = None     # on ``dst`` only: the whole-band mix before FIR2
    mix: Optional[np.ndarray] = None         #                  after FIR2
    pcm: Optional[np.ndarray] = None         #                  s16


class ShardedReceiver:
    """``engine_factory(sub_plan) -> engine`` builds this rank's engine; the default is the CUDA engine with
    ``partial_mix=True`` on ``device``.  (The CPU tests inject an oracle-backed stand-in: the product has no
    CPU engine.)  An engine needs ``process(raw, fmt) -> BlockOut`` and ``finish_mix(parts) -> (mix_pre, mix, pcm)``."""

    def __init__(self, plan: ChannelPlan, fmt: int = FMT_F32, device: Optional[int] = None, src: int = 0, dst: int = 0,
                 engine_factory: Optional[Callable[[ChannelPlan], object]] = None, group=None, **engine_kw):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.src, self.dst, self.fmt = src, dst, fmt
        self.plan = plan
        self.lo, self.hi = channel_range(len(plan.offsets), self.world, self.rank)
        if self.hi == self.lo:
            raise ValueError(f"rank {self.rank} would own no channel: {len(plan.offsets)} channels over {self.world} ranks")
        sub = shard_plan(plan, self.world, self.rank)
        self.on_gpu = dist.get_backend(group) == "nccl"
        if engine_factory is None:
            from .engine import Engine
            dev = device if device is not None else (torch.cuda.current_device() if self.on_gpu else 0)
            self.device = dev
            self.engine = Engine(sub, device=dev, partial_mix=True, **engine_kw)
        else:
            self.device = device
            self.engine = engine_factory(sub)
        self.block_len = plan.rate.block_len
        n_elem = 2 * self.block_len
        tdt = {FMT_U8: torch.uint8, FMT_S16: torch.int16, FMT_F32: torch.float32}[fmt]
        comm_dev = torch.device("cuda", self.device) if self.on_gpu else torch.device("cpu")
        self._blk = torch.empty(n_elem, dtype=tdt, device=comm_dev)           # the block, wire format
        self._part = torch.empty(2 * CHUNK, dtype=torch.float32, device=comm_dev)
        self._parts = ([torch.empty(2 * CHUNK, dtype=torch.float32, device=comm_dev) for _ in range(self.world)]
                       if self.rank == dst else None)

    def process(self, raw: Optional[np.ndarray]) -> ShardOut:
        """``raw``: the block in the wire format on rank ``src`` (ignored elsewhere)."""
        torch, dist = self.torch, self.dist
        if self.rank == self.src:
            a = np.ascontiguousarray(raw)
            if a.dtype == np.complex64:
                a = a.view(np.float32)
            if a.dtype != _WIRE_DTYPE[self.fmt] or a.size != 2 * self.block_len:
                raise ValueError("block does not match the receiver's wire format / block length")
            self._blk.copy_(torch.from_numpy(a), non_blocking=False)
        dist.broadcast(self._blk, src=self.src, group=self.group)
        if self.on_gpu:
            torch.cuda.current_stream().synchronize()                    # the engine reads it on its own stream
            self.engine.push_device(self._blk.data_ptr(), self.fmt, self.block_len)
            local = self.engine.outputs()
        else:
            local = self.engine.process(self._blk.numpy(), self.fmt)
        self._part.copy_(torch.from_numpy(np.ascontiguousarray(local.mix_pre, np.float32)))
        dist.gather(self._part, self._parts, dst=self.dst, group=self.group)
        out = ShardOut(self.lo, self.hi, local)
        if self.rank == self.dst:
            parts = np.stack([p.cpu().numpy() for p in self._parts])
            out.mix_pre, out.mix, out.pcm = self.engine.finish_mix(parts)
        return out

    def close(self) -> None:
        if hasattr(self.engine, "close"):
            self.engine.close()

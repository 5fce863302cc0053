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


class _DevArray:
    """A device pointer the engine owns, dressed for ``torch.as_tensor`` (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, n: int, typestr: str = "<f4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class ShardedReceiver:
    """``engine_factory(sub_plan) -> engine`` builds this rank's engine; the default is the CUDA engine with
    ``partial_mix=True`` on ``device``.  (The CPU tests inject an oracle-backed stand-in: the product has no
    CPU engine.)  An engine needs ``process(raw, fmt) -> BlockOut`` and ``finish_mix(parts) -> (mix_pre, mix, pcm)``.

    On GPUs (``nccl`` backend) the receiver is a pipeline without host synchronisation on the data path:

    * ``submit(raw)``: rank ``src`` copies the block (pinned host memory, wire format) to its GPU on a comm stream and
      one ``ncclBroadcast`` hands it to every rank (``ingest="broadcast"``); or every rank copies its own 1/world slice
      over its own host link and one ``ncclAllGather`` assembles the block (``ingest="allgather"``, for replay faster
      than one host link allows).  The engine's ingest is ordered behind the collective on the device
      (``sdrx_engine_wait_stream``), the block is pushed, the per-rank results (squelch, SNR, sparse audio) are fetched
      asynchronously, and the partial ``mix_pre`` of the block -- straight from the engine's device buffers -- goes
      through one small ``ncclAllGather`` on a second communicator, behind the block's boundary B; ``dst`` finishes the
      mix (FIR2, s16) on the same stream (``sdrx_engine_finish_mix_async``).
    * ``collect()``: the oldest outstanding block's results, ``depth`` blocks may be in flight.
    * ``process(raw)`` = ``submit`` + ``collect`` (the synchronous form the parity tests use)."""

    def __init__(self, plan: ChannelPlan, fmt: int = FMT_F32, device: Optional[int] = None, src: int = 0, dst: int = 0,
                 engine_factory: Optional[Callable[[ChannelPlan], object]] = None, group=None, ingest: str = "broadcast",
                 depth: int = 4, want: Tuple[str, ...] = ("audio_rows", "audio_chan", "n_audio", "sql_state", "snr"),
                 **engine_kw):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.src, self.dst, self.fmt = src, dst, fmt
        self.plan = plan
        self.lo, self.hi = channel_range(len(plan.offsets), self.world, self.rank)
        if self.hi == self.lo:
            raise ValueError(f"rank {self.rank} would own no channel: {len(plan.offsets)} channels over {self.world} ranks")
        if ingest not in ("broadcast", "allgather"):
            raise ValueError("ingest is 'broadcast' or 'allgather'")
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
        self.ingest, self.depth = ingest, max(1, min(int(depth), 4))
        n_elem = 2 * self.block_len
        tdt = {FMT_U8: torch.uint8, FMT_S16: torch.int16, FMT_F32: torch.float32}[fmt]
        self._tdt = tdt
        comm_dev = torch.device("cuda", self.device) if self.on_gpu else torch.device("cpu")
        if not self.on_gpu:
            self._blk = torch.empty(n_elem, dtype=tdt, device=comm_dev)           # the block, wire format
            self._part = torch.empty(2 * CHUNK, dtype=torch.float32, device=comm_dev)
            self._parts = ([torch.empty(2 * CHUNK, dtype=torch.float32, device=comm_dev) for _ in range(self.world)]
                           if self.rank == dst else None)
            return
        # ---- GPU pipeline state
        if n_elem % self.world:
            raise ValueError("block length must divide by the world size for ingest='allgather'")
        self._lib = self.engine.lib
        self._h = self.engine.h
        self._blk = [torch.empty(n_elem, dtype=tdt, device=comm_dev) for _ in range(self.depth)]
        self._parts = [torch.empty(self.world * 2 * CHUNK, dtype=torch.float32, device=comm_dev) for _ in range(self.depth)]
        self._s_in = torch.cuda.Stream(device=comm_dev)         # block in: H2D + broadcast / all-gather
        self._s_mix = torch.cuda.Stream(device=comm_dev)        # partial mixes out: all-gather + finish on dst
        # the mix gather waits for boundary B of its block; on its own communicator it cannot hold up the next block's
        # broadcast inside NCCL's stream
        self._mix_group = dist.new_group(ranks=list(range(self.world)), backend="nccl") if group is None else group
        self._osets = [self.engine.make_output_buffers(pinned=True, want=want) for _ in range(self.depth)]
        from .engine import PinnedBuffer
        self._fin = [(PinnedBuffer(8 * CHUNK), PinnedBuffer(8 * CHUNK), PinnedBuffer(4 * CHUNK)) for _ in range(self.depth)] \
            if self.rank == dst else None
        self._stage = [PinnedBuffer(n_elem * self._blk[0].element_size()) for _ in range(self.depth)]
        self._done = [torch.cuda.Event() for _ in range(self.depth)]
        self._mixpre_view = {}
        self._n_sub = self._n_col = 0

    # ------------------------------------------------------------------------------------------ GPU pipeline
    def _mixpre_tensor(self):
        """The engine's device-side mix_pre of the block just pushed, as a torch tensor (no copy)."""
        import ctypes as C
        from . import _lib
        o = _lib.BlockOutputs()
        _lib.check(self._lib.sdrx_engine_device_outputs(self._h, 0, C.byref(o)))
        ptr = C.cast(o.mix_pre, C.c_void_p).value
        t = self._mixpre_view.get(ptr)
        if t is None:
            t = self.torch.as_tensor(_DevArray(ptr, 2 * CHUNK), device=self.torch.device("cuda", self.device))
            self._mixpre_view[ptr] = t
        return t

    def submit(self, raw: Optional[np.ndarray] = None, device_block=None) -> None:
        """Queue one block.  ``raw``: the block in the wire format -- on rank ``src`` for ``ingest="broadcast"``, on every
        rank (each reads only its own slice) for ``ingest="allgather"``; a :class:`PinnedBuffer` view is copied straight
        from where it lies.  ``device_block``: a torch tensor already on this rank's GPU instead of ``raw``."""
        import ctypes as C
        from . import _lib
        torch, dist = self.torch, self.dist
        if self._n_sub - self._n_col >= self.depth:
            raise RuntimeError("ShardedReceiver: collect() before submitting more than `depth` blocks")
        k = self._n_sub % self.depth
        blk = self._blk[k]

        def host_tensor(a, lo, hi):
            a = np.ascontiguousarray(a)
            if a.dtype == np.complex64:
                a = a.view(np.float32)
            if a.dtype != _WIRE_DTYPE[self.fmt] or a.size != 2 * self.block_len:
                raise ValueError("block does not match the receiver's wire format / block length")
            t = torch.from_numpy(a[lo:hi])
            if not t.is_pinned():                                    # pageable caller: through this slot's pinned stage
                st = torch.from_numpy(self._stage[k].view(_WIRE_DTYPE[self.fmt], 2 * self.block_len)[lo:hi])
                st.copy_(t)
                t = st
            return t

        with torch.cuda.stream(self._s_in):
            if self.ingest == "broadcast":
                if self.rank == self.src:
                    if device_block is not None:
                        blk.copy_(device_block.view(self._tdt).reshape(-1), non_blocking=True)
                    else:
                        blk.copy_(host_tensor(raw, 0, 2 * self.block_len), non_blocking=True)
                dist.broadcast(blk, src=self.src, group=self.group)
            else:
                n = 2 * self.block_len // self.world
                mine = blk[self.rank * n:(self.rank + 1) * n]
                if device_block is not None:
                    mine.copy_(device_block.view(self._tdt).reshape(-1)[self.rank * n:(self.rank + 1) * n], non_blocking=True)
                else:
                    mine.copy_(host_tensor(raw, self.rank * n, (self.rank + 1) * n), non_blocking=True)
                dist.all_gather_into_tensor(blk, mine, group=self.group)
        _lib.check(self._lib.sdrx_engine_wait_stream(self._h, C.c_void_p(self._s_in.cuda_stream)))
        self.engine.push_device(blk.data_ptr(), self.fmt, self.block_len)
        # the comm buffer of this slot is free again once the engine has copied the block into its raw ring
        _lib.check(self._lib.sdrx_engine_stream_wait_ingest(self._h, C.c_void_p(self._s_in.cuda_stream)))
        self.engine.fetch_outputs(self._osets[k][0])
        _lib.check(self._lib.sdrx_engine_stream_wait_outputs(self._h, C.c_void_p(self._s_mix.cuda_stream)))
        part = self._mixpre_tensor()
        with torch.cuda.stream(self._s_mix):
            dist.all_gather_into_tensor(self._parts[k], part, group=self._mix_group)
            if self.rank == self.dst:
                f = self._fin[k]
                _lib.check(self._lib.sdrx_engine_finish_mix_async(
                    self._h, C.c_void_p(self._parts[k].data_ptr()), self.world, C.c_void_p(self._s_mix.cuda_stream),
                    C.cast(f[0].ptr, _lib.c_float_p), C.cast(f[1].ptr, _lib.c_float_p), C.cast(f[2].ptr, _lib.c_i16_p)))
            self._done[k].record(self._s_mix)
        self._n_sub += 1

    def collect(self) -> ShardOut:
        """Results of the oldest block in flight (blocks until they are in host memory).  ``local`` holds this rank's
        output arrays (views of pinned buffers that the slot's next block overwrites: copy what must be kept)."""
        if self._n_col >= self._n_sub:
            raise RuntimeError("ShardedReceiver: nothing in flight")
        k = self._n_col % self.depth
        self.engine.wait_outputs(self._n_sub - 1 - self._n_col)
        self._done[k].synchronize()
        out = ShardOut(self.lo, self.hi, self._osets[k][1])
        if self.rank == self.dst:
            f = self._fin[k]
            out.mix_pre, out.mix, out.pcm = f[0].view(np.float32, 2 * CHUNK), f[1].view(np.float32, 2 * CHUNK), f[2].view(np.int16, 2 * CHUNK)
        self._n_col += 1
        return out

    def process(self, raw: Optional[np.ndarray]) -> ShardOut:
        """``raw``: the block in the wire format on rank ``src`` (ignored elsewhere)."""
        torch, dist = self.torch, self.dist
        if self.on_gpu:
            self.submit(raw)
            return self.collect()
        if self.rank == self.src:
            a = np.ascontiguousarray(raw)
            if a.dtype == np.complex64:
                a = a.view(np.float32)
            if a.dtype != _WIRE_DTYPE[self.fmt] or a.size != 2 * self.block_len:
                raise ValueError("block does not match the receiver's wire format / block length")
            self._blk.copy_(torch.from_numpy(a), non_blocking=False)
        dist.broadcast(self._blk, src=self.src, group=self.group)
        local = self.engine.process(self._blk.numpy(), self.fmt)
        self._part.copy_(torch.from_numpy(np.ascontiguousarray(local.mix_pre, np.float32)))
        dist.gather(self._part, self._parts, dst=self.dst, group=self.group)
        out = ShardOut(self.lo, self.hi, local)
        if self.rank == self.dst:
            parts = np.stack([p.cpu().numpy() for p in self._parts])
            out.mix_pre, out.mix, out.pcm = self.engine.finish_mix(parts)
        return out

    def close(self) -> None:
        if self.on_gpu:
            self.torch.cuda.synchronize()
        if hasattr(self.engine, "close"):
            self.engine.close()

"""World-size-2 gloo tests (CPU) of the channel-sharded receiver's host logic (sdrx_b200/sharded.py):
channel partition, wire-format block broadcast from the source rank, per-rank partial mixes gathered and
finished in rank order.  The per-rank engine is an oracle-backed stand-in injected by the test (the product
has no CPU engine); the GPU version of the same check is tests/test_gpu_sharded.py."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from sdrx_b200.sharded import channel_range, shard_plan
from tests import scenarios as sc


def test_channel_range_is_a_balanced_contiguous_partition():
    for n, w in ((960, 8), (960, 7), (8, 2), (10, 4), (5, 5), (128, 3)):
        r = [channel_range(n, w, k) for k in range(w)]
        assert r[0][0] == 0 and r[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
        sizes = [hi - lo for lo, hi in r]
        assert max(sizes) - min(sizes) <= 1


def test_shard_plan_keeps_global_pan_positions():
    plan = sc.config2().plan()
    a, b = shard_plan(plan, 2, 0), shard_plan(plan, 2, 1)
    assert a.offsets + b.offsets == plan.offsets
    assert a.pos + b.pos == plan.pos == [-2, -2, -1, -1, 1, 1, 2, 2]       # sdrx.cpp:1024-1044 on the global list


class _OracleShardEngine:
    """Stand-in with the engine surface ShardedReceiver needs, backed by the CPU oracle (tests only)."""

    def __init__(self, sub, scn):
        from oracle import bindings as ob
        from sdrx_b200 import config as cf
        from tests.harness import product_tables
        self.rx = ob.OracleRx(product_tables(scn.rate), sub.offsets, sub.sql_levels, sub.pos, scn.use_lf_agc, scn.lf_gain_db)
        self.fir2 = ob.OracleFIR(2, np.ascontiguousarray(cf.tables()["coeff_bp4am_channel"], np.float32))

    def process(self, raw, fmt):
        return self.rx.process(np.array(raw, copy=True), fmt)

    def finish_mix(self, parts):
        acc = parts[0].astype(np.float32).copy()
        for p in parts[1:]:
            acc = (acc + p.astype(np.float32)).astype(np.float32)
        mix = self.fir2.filter(acc)
        pcm = np.where(mix > 1.0, 32767, np.where(mix < -1.0, -32767, (mix * np.float32(32767.0)).astype(np.int16))).astype(np.int16)
        return acc, mix, pcm


def _worker(rank, world, port, n_blocks, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sdrx_b200.sharded import ShardedReceiver
        scn = sc.config2(n_blocks=n_blocks)
        rx = ShardedReceiver(scn.plan(), fmt=scn.fmt, engine_factory=lambda sub: _OracleShardEngine(sub, scn))
        src = scn.source()
        outs = []
        for b in range(n_blocks):
            o = rx.process(src.block(b) if rank == 0 else None)       # only the source rank has the block
            outs.append((o.lo, o.hi, o.local.audio, o.local.sql, o.local.ch_iq, o.mix_pre, o.mix, o.pcm))
        q.put((rank, outs))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(600)
def test_two_rank_channel_sharding_matches_the_single_receiver():
    from tests.harness import AUDIO_TOL, rel_rms, run_cpu
    n_blocks, world = 6, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_blocks, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=500) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    scn = sc.config2(n_blocks=n_blocks)
    want = run_cpu(scn, "oracle")
    assert any(w.sql.any() for w in want)
    for b in range(n_blocks):
        audio = np.concatenate([got[r][b][2] for r in range(world)])
        sql = np.concatenate([got[r][b][3] for r in range(world)])
        iq = np.concatenate([got[r][b][4] for r in range(world)])
        assert (got[0][b][0], got[0][b][1], got[1][b][0], got[1][b][1]) == (0, 4, 4, 8)
        # everything per channel is identical to the unsharded run
        assert np.array_equal(audio, want[b].audio) and np.array_equal(sql, want[b].sql) and np.array_equal(iq, want[b].ch_iq)
        # the mix differs only by the association of the adds between ranks
        assert got[1][b][5] is None and got[1][b][7] is None
        assert rel_rms(got[0][b][6], want[b].mix) < AUDIO_TOL or not np.any(want[b].mix)
        assert np.max(np.abs(got[0][b][7].astype(np.int32) - want[b].pcm.astype(np.int32))) <= 1

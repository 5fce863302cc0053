"""GPU check of the channel-sharded pieces of the C ABI (SDRX_FLAG_PARTIAL_MIX, sdrx_engine_finish_mix): two
engines that each own half of the channels of one stream, on one GPU, against the unsharded engine."""
import numpy as np
import pytest

from sdrx_b200.sharded import shard_plan
from tests import scenarios as sc
from tests.harness import AUDIO_TOL, rel_rms

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("force_generic", [True, False])
def test_two_channel_shards_reproduce_the_unsharded_engine(force_generic):
    from sdrx_b200 import Engine
    scn = sc.config2(n_blocks=6)
    src = scn.source()
    blocks = [src.block(b) for b in range(scn.n_blocks)]
    plan = scn.plan()
    whole = Engine(plan, force_generic=force_generic)
    parts = [Engine(shard_plan(plan, 2, r), force_generic=force_generic, partial_mix=True) for r in range(2)]
    one = Engine(plan, force_generic=force_generic, partial_mix=True)      # a single "shard": must be bit-exact
    for blk in blocks:
        w = whole.process(blk)
        outs = [p.process(blk) for p in parts]
        assert np.array_equal(np.concatenate([o.audio for o in outs]), w.audio)
        assert np.array_equal(np.concatenate([o.sql for o in outs]), w.sql)
        assert np.array_equal(np.concatenate([o.ch_iq for o in outs]), w.ch_iq)
        mix_pre, mix, pcm = parts[0].finish_mix(np.stack([o.mix_pre for o in outs]))
        assert rel_rms(mix, w.mix) < AUDIO_TOL or not np.any(w.mix)
        assert np.max(np.abs(pcm.astype(np.int32) - w.pcm.astype(np.int32))) <= 1
        o1 = one.process(blk)
        m1 = one.finish_mix(o1.mix_pre[None, :])
        assert np.array_equal(m1[0], w.mix_pre) and np.array_equal(m1[1], w.mix) and np.array_equal(m1[2], w.pcm)
    for e in [whole, one] + parts:
        e.close()


def test_channel_sharded_receiver_over_nccl():
    """Row e2: the pipelined ShardedReceiver with the nccl backend on two GPUs -- NCCL broadcast (and the
    all-gather ingest) of every block, NCCL gather of the partial mixes, mix finished on rank 0 -- against a single-GPU
    engine on the same blocks.  Skips below two devices."""
    import json
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29571", os.path.join(here, "nccl_sharded_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("NCCL_SHARDED ")][-1]
    res = json.loads(line[len("NCCL_SHARDED "):])
    print(line)
    for ingest in ("broadcast", "allgather"):
        assert res[ingest]["squelch_equal"]
        assert res[ingest]["worst_audio_rel_rms"] < 1e-4 and res[ingest]["worst_mix_rel_rms"] < 1e-4
        assert res[ingest]["worst_pcm_lsb"] <= 2

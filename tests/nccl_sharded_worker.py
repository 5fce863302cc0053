"""Worker of tests/test_gpu_sharded.py::test_channel_sharded_receiver_over_nccl (one process per GPU, torchrun):
the pipelined ShardedReceiver (NCCL broadcast / all-gather of the block, NCCL gather of the partial mixes) against a
single-GPU engine on the same blocks.  Rank 0 prints one JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from sdrx_b200 import Engine
    from sdrx_b200.sharded import ShardedReceiver
    from tests import scenarios as sc
    from tests.harness import rel_rms
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    scn = sc.paired_channels()
    scn.n_blocks = 9
    src = scn.source()
    blocks = [src.block(b) for b in range(scn.n_blocks)]
    plan = scn.plan()
    res = {}
    for ingest in ("broadcast", "allgather"):
        rx = ShardedReceiver(plan, device=local, ingest=ingest, depth=3)
        outs = []

        def keep(o):
            a = o.local
            n = int(a["n_audio"][0])
            audio = np.zeros((rx.hi - rx.lo, 512), np.float32)
            audio[a["audio_chan"][:n]] = a["audio_rows"][:n * 512].reshape(n, 512)
            outs.append(dict(audio=audio, sql=a["sql_state"].copy(), snr=a["snr"].copy(),
                             pcm=None if o.pcm is None else o.pcm.copy(), mix=None if o.mix is None else o.mix.copy()))
        for i, b in enumerate(blocks):                       # pipelined: up to three blocks in flight
            rx.submit(b if (rank == 0 or ingest == "allgather") else None)
            if i >= 2:
                keep(rx.collect())
        keep(rx.collect())
        keep(rx.collect())
        rx.close()
        gathered = [None] * world
        dist.gather_object(outs, gathered if rank == 0 else None, dst=0)
        if rank == 0:
            eng = Engine(plan, device=local)
            ref = [eng.process(b) for b in blocks]
            eng.close()
            ok_sql, worst_audio, worst_mix, worst_pcm = True, 0.0, 0.0, 0
            for b, r in enumerate(ref):
                audio = np.concatenate([gathered[k][b]["audio"] for k in range(world)])
                sql = np.concatenate([gathered[k][b]["sql"] for k in range(world)])
                ok_sql = ok_sql and bool(np.array_equal(sql, r.sql))
                assert np.array_equal(audio == 0.0, r.audio == 0.0), "gating differs"
                for c in range(audio.shape[0]):
                    if np.any(r.audio[c] != 0.0):
                        worst_audio = max(worst_audio, rel_rms(audio[c], r.audio[c]))
                if np.any(r.mix != 0.0):
                    worst_mix = max(worst_mix, rel_rms(gathered[0][b]["mix"], r.mix))
                worst_pcm = max(worst_pcm, int(np.max(np.abs(gathered[0][b]["pcm"].astype(np.int32) - r.pcm.astype(np.int32)))))
            res[ingest] = dict(squelch_equal=ok_sql, worst_audio_rel_rms=worst_audio, worst_mix_rel_rms=worst_mix,
                               worst_pcm_lsb=worst_pcm, blocks=len(ref))
    if rank == 0:
        print("NCCL_SHARDED " + json.dumps(dict(world=world, **res)))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

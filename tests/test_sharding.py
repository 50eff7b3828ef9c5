"""CPU: the N>1 host logic (partition + ordered merge) under torch.distributed/gloo, world size 2.
The per-channel decoder is a stand-in here (the oracle) -- what is under test is the sharding."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partitions():
    from msk144cudecoder_b200.sharding import channels_of_rank
    for n in (0, 1, 7, 8, 1024):
        for w in (1, 2, 4, 8):
            parts = [channels_of_rank(n, r, w) for r in range(w)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1


def test_batch_form_maps_local_channels_back():
    from msk144cudecoder_b200.sharding import decode_sharded_batch
    seen = []

    def decode_channels(chans):
        seen.append(list(chans))
        return [dict(channel=i, block=0, text=f"c{c}") for i, c in enumerate(chans)]

    out = decode_sharded_batch(decode_channels, 7, 1, 3, gather=False)
    assert seen == [[3, 4]] and [(l["channel"], l["text"]) for l in out] == [(3, "c3"), (4, "c4")]


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from msk144cudecoder_b200 import synth
    from msk144cudecoder_b200.sharding import decode_sharded
    from oracle import orc
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    cfg = orc.config(width=60, step=2, depth=2, thr=1, nthreads=2)

    def decode_channel(c):
        x, _ = synth.cfg5_channel(c, seconds=2.0)
        return orc.decode_stream(x, cfg)[0]

    lines = decode_sharded(decode_channel, 3, rank, world)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, [(l["channel"], l["block"], l["text"]) for l in lines]))


def test_two_rank_gather_equals_single_process():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    sys.path.insert(0, ROOT)
    from msk144cudecoder_b200 import synth
    from msk144cudecoder_b200.sharding import decode_sharded
    from oracle import orc
    cfg = orc.config(width=60, step=2, depth=2, thr=1, nthreads=2)
    single = decode_sharded(lambda c: orc.decode_stream(synth.cfg5_channel(c, seconds=2.0)[0], cfg)[0], 3, 0, 1)
    assert got[0] == [(l["channel"], l["block"], l["text"]) for l in single]
    assert set(got[1]) <= set(got[0])           # rank 1 keeps only its own share

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


def _fake_stream(nblocks=40):
    """What the GPU stages of one channel would return, made on the CPU: per-block sub-block powers with a noise-floor
    step, and LDPC successes whose TEXT depends on stream order (a type-4 message teaches a hash that later messages
    carry)."""
    import msk144cudecoder_b200 as m
    from msk144cudecoder_b200 import synth
    rng = np.random.default_rng(9)
    powers = (5184.0 / 8 * (1.0 + 0.05 * rng.standard_normal((nblocks, 8)))).astype(np.float32)
    powers[nblocks // 2:] *= 3.0
    powers[5] *= 40.0
    powers[30] *= 25.0
    say = {5: "CQ PJ4/K1ABC", 6: "CQ PJ4/K1ABC", 12: "K1ABC W9XYZ EN37", 30: "<PJ4/K1ABC> W9XYZ RR73", 37: "W9XYZ <PJ4/K1ABC> RRR"}
    res = np.zeros(2 * len(say), m.RESULT_DTYPE)
    for i, (b, text) in enumerate(sorted(say.items())):
        for j in range(2):          # two candidates of the block decode the same payload
            r = res[2 * i + j]
            r["block_index"], r["item_index"], r["pattern_idx"], r["num_avg"] = b, 100 + 8 * j, j, j + 1
            r["f0"], r["nbadsync"] = 1500.0 + j, 1 - j
            r["message"][:77] = synth.pack77(text)
    return powers, res


def _time_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import msk144cudecoder_b200 as m
    from msk144cudecoder_b200.sharding import decode_time_sharded
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    powers, res = _fake_stream()

    def decode_chunk(b0, b1):       # stand-in for Decoder.decode_chunk_host on the samples of blocks [b0, b1)
        sel = (res["block_index"] >= b0) & (res["block_index"] < b1)
        return res[sel].copy(), powers[b0:b1].copy()

    lines = decode_time_sharded(decode_chunk, m.replay_channel, powers.shape[0], rank, world)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, [(l["block"], l["text"], l["snr"], l["num_avg"], l["nbadsync"]) for l in lines]))


def test_two_rank_time_pieces_replayed_in_order():
    """One channel cut in time over two ranks: rank 0 gathers (powers, LDPC successes) and replays SNR and text once
    in block order; the lines equal the undivided pass, including the hashed call sign learnt in rank 0's half and
    used in rank 1's."""
    import torch.multiprocessing as mp
    import msk144cudecoder_b200 as m
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_time_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    powers, res = _fake_stream()
    whole = [(l["block"], l["text"], l["snr"], l["num_avg"], l["nbadsync"]) for l in m.replay_channel(powers, res)]
    assert got[0] == whole and got[1] == []
    assert [t for _, t, *_ in whole] == ["CQ PJ4/K1ABC", "CQ PJ4/K1ABC", "K1ABC W9XYZ EN37", "<PJ4/K1ABC> W9XYZ RR73",
                                         "W9XYZ <PJ4/K1ABC> RRR"]
    assert len({snr for _, _, snr, *_ in whole}) > 1
    # rank 1's half replayed on its own would not know the call sign
    late = m.replay_channel(powers[20:], res[res["block_index"] >= 20], first_block=20)
    assert [l["text"] for l in late] == ["<...> W9XYZ RR73", "W9XYZ <...> RRR"]
    # out-of-order input is refused, not silently mis-replayed
    with pytest.raises(m.Msk144Error, match="not in (block|item) order"):
        m.replay_channel(powers, res[::-1].copy())

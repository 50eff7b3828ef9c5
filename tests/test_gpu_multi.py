"""GPU: more than one decoder at work -- the in-process pool (msk144_pool_*, one host thread per member) and the
multi-process sharding (torch.distributed, one rank per GPU) must both return exactly what ONE decoder returns.
On a one-GPU box the members / ranks share GPU 0; with two or more GPUs they are spread (nccl for the ranks)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KW = dict(width=100, step=2, depth=3, thr=2)
NCHAN, SECONDS = 5, 3.0


@pytest.fixture(scope="module")
def gpu(product):
    import torch
    assert torch.cuda.is_available()
    return product


def channels(synth, n=NCHAN):
    return np.stack([synth.cfg5_channel(c, seconds=SECONDS)[0] for c in range(n)])


def strip(lines):
    return [(l["channel"], l["block"], l["text"], l["snr"], l["f0"], l["num_avg"], l["nbadsync"], l["pattern_idx"], l["payload"])
            for l in lines]


@pytest.mark.gpu
def test_pool_batch_equals_single_decoder(gpu, synth):
    import torch
    x = channels(synth)
    with gpu.Decoder(compat_cache=True, **KW) as d:
        single, sres = d.decode_host(x, nchan=NCHAN)
    assert len(single) > 0 and len({l["channel"] for l in single}) > 1
    ndev = torch.cuda.device_count()
    for devices in ([0], [0, 0, 0], [i % ndev for i in range(2)], [0] * 7):   # 7 members > 5 channels: two stay idle
        with gpu.Pool(devices, compat_cache=True, **KW) as p:
            pooled, pres = p.decode_host(x, nchan=NCHAN, result_cap=1 << 16)
        assert strip(pooled) == strip(single), devices
        assert pres.tobytes() == sres.tobytes(), devices


@pytest.mark.gpu
def test_pool_streaming_equals_single_decoder(gpu, synth):
    x = channels(synth)
    il = np.ascontiguousarray(x.T)          # [t][c]: one sample of every channel per time step
    raw = il.tobytes()
    frame = 2 * NCHAN
    hop = 2592 * frame
    with gpu.Decoder(compat_cache=True, **KW) as d:
        single = []
        for off in range(0, len(raw), hop):
            d.push(raw[off:off + hop], nchan=NCHAN)
            single += d.poll()
    with gpu.Pool([0, 0], compat_cache=True, **KW) as p:
        pooled = []
        for off in range(0, len(raw), hop):
            p.push(raw[off:off + hop], NCHAN)
            pooled += p.poll()
        with pytest.raises(gpu.Msk144Error):
            p.push(raw[:frame + 1], NCHAN)      # not a whole number of time steps
        p.reset_stream()
        p.push(raw[:2 * hop], NCHAN)            # usable again after a reset
    assert len(single) > 0 and strip(pooled) == strip(single)


@pytest.mark.gpu
@pytest.mark.parametrize("members,flags", [(NCHAN, 0), (NCHAN + 2, 0), (NCHAN, 2)])
def test_pool_streaming_one_channel_per_member(gpu, synth, members, flags):
    """As many members as channels (and more: the extra ones own nothing): every member streams ONE column.  It runs
    the interleaved form like the single decoder does, so MSK144_FLAG_SKIP_WAV_HEADER (flags=2, which only the
    one-channel msk144_push honours) changes nothing."""
    x = channels(synth)
    raw = np.ascontiguousarray(x.T).tobytes()
    hop = 2592 * 2 * NCHAN
    with gpu.Decoder(compat_cache=True, flags=flags, **KW) as d:
        single = []
        for off in range(0, len(raw), 3 * hop):
            d.push(raw[off:off + 3 * hop], nchan=NCHAN)
            single += d.poll()
    with gpu.Pool([0] * members, compat_cache=True, flags=flags, **KW) as p:
        pooled = []
        for off in range(0, len(raw), 2 * hop):     # chunked differently from the single decoder on purpose
            p.push(raw[off:off + 2 * hop], NCHAN)
            pooled += p.poll()
    assert len(single) > 0 and strip(pooled) == strip(single)


@pytest.mark.gpu
def test_pool_streaming_member_failure_keeps_lines_and_asks_for_reset(gpu, synth):
    """One member fails in mid-stream: the push reports it, the lines the other members produced from that push are
    still delivered, further pushes are refused until msk144_pool_reset_stream, and after the reset the pool works."""
    import ctypes as C
    x = channels(synth)
    raw = np.ascontiguousarray(x.T).tobytes()
    frame = 2 * NCHAN
    L = gpu.lib()
    with gpu.Pool([0, 0], compat_cache=True, **KW) as p:
        p.push(raw, NCHAN)
        healthy = p.poll()
        assert len(healthy) > 0
        p.reset_stream()
        # member 1 (channels 3, 4) is put out of step behind the pool's back: its stream now expects 3 channels
        m1 = L.msk144_pool_member(p._h, 1)
        assert m1 and L.msk144_push_interleaved(m1, None, 0, 3) == 0
        with pytest.raises(gpu.Msk144Error, match="device 0: channel count changed"):
            p.push(raw, NCHAN)
        kept = p.poll()                              # member 0's lines of the failed push
        assert strip(kept) == strip([l for l in healthy if l["channel"] < 3]) and len(kept) > 0
        with pytest.raises(gpu.Msk144Error, match="earlier push failed"):
            p.push(raw[:frame], NCHAN)
        p.reset_stream()
        p.push(raw, NCHAN)
        assert strip(p.poll()) == strip(healthy)


@pytest.mark.gpu
def test_cli_devices_equals_single_device(gpu, synth):
    import subprocess
    x = channels(synth)
    raw = np.ascontiguousarray(x.T).tobytes()
    base = [gpu.CLI_PATH, "--search-width=100", "--search-step=2", "--scan-depth=3", "--nbadsync-threshold=2", f"--channels={NCHAN}"]
    one = subprocess.run(base, input=raw, capture_output=True)
    two = subprocess.run(base + ["--devices=0,0"], input=raw, capture_output=True)
    assert one.returncode == 0 and two.returncode == 0, two.stderr.decode()[-300:]
    mask = lambda b: [l.split("date=")[0] + l.split("date=")[1][14:] if "date=" in l else l for l in b.decode().splitlines()]
    assert len(mask(one.stdout)) > 1 and mask(one.stdout) == mask(two.stdout)
    assert b"GPUs: 2" in two.stderr


def _rank(rank, world, port, backend, ndev, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import msk144cudecoder_b200 as m
    from msk144cudecoder_b200 import synth
    from msk144cudecoder_b200.sharding import decode_sharded_batch
    dev = rank % ndev
    torch.cuda.set_device(dev)
    dist.init_process_group(backend, init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world,
                            **({"device_id": torch.device("cuda", dev)} if backend == "nccl" else {}))
    with m.Decoder(device=dev, compat_cache=True, **KW) as d:
        def decode_channels(chans):
            x = np.stack([synth.cfg5_channel(c, seconds=SECONDS)[0] for c in chans])
            return d.decode_host(x, nchan=len(chans))[0]
        lines = decode_sharded_batch(decode_channels, NCHAN, rank, world)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, strip(lines)))


@pytest.mark.gpu
def test_two_ranks_gather_equals_single_gpu(gpu, synth):
    """sharding.decode_sharded_batch with the CUDA decoder on every rank: rank 0's merged lines == one decoder's."""
    import torch
    import torch.multiprocessing as mp
    ndev = torch.cuda.device_count()
    backend = "nccl" if ndev >= 2 else "gloo"
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank, args=(r, 2, port, backend, ndev, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=600) for _ in range(2))
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    with gpu.Decoder(compat_cache=True, **KW) as d:
        single, _ = d.decode_host(channels(synth), nchan=NCHAN)
    key = lambda ls: sorted(ls, key=lambda t: (t[0], t[1], t[2]))
    assert got[0] == key(strip(single)) and len(got[0]) > 0
    assert set(got[1]) < set(got[0])           # rank 1 keeps only its own share


@pytest.mark.gpu
def test_small_output_arrays_do_not_cost_a_second_decode(gpu, synth):
    """MSK144_EOVERFLOW reports the sizes needed; msk144_last_output / msk144_pool_last_output then hand over the
    output the decoder kept -- the kernels run once (kernel_launches of the stage times), and the lines and results
    are the ones a call with large arrays returns."""
    x = channels(synth)
    with gpu.Decoder(compat_cache=True, **KW) as d:
        big, bres = d.decode_host(x, nchan=NCHAN, line_cap=4096, result_cap=1 << 16)
        launches_one = d.stage_times()["kernel_launches"]
        small, sres = d.decode_host(x, nchan=NCHAN, line_cap=1, result_cap=1)
        launches_two = d.stage_times()["kernel_launches"]
    assert len(big) > 1 and len(bres) > 1
    assert strip(small) == strip(big) and sres.tobytes() == bres.tobytes()
    assert launches_two - launches_one == launches_one - 1, (launches_one, launches_two)   # minus the one-off mixer table
    with gpu.Pool([0, 0], compat_cache=True, **KW) as p:
        pooled, pres = p.decode_host(x, nchan=NCHAN, line_cap=1, result_cap=1)
    assert strip(pooled) == strip(big) and pres.tobytes() == bres.tobytes()


def hashed_call_stream(synth, seconds=24.0):
    """One channel whose text depends on stream order: a type-4 message early teaches the hash of PJ4/K1ABC, messages
    that carry only that hash follow late (they print <PJ4/K1ABC> after it, <...> before), a strong ping moves the SNR
    noise floor; the rest is cfg2-like noise with weak pings."""
    n = int(seconds * 12000)
    P = synth.Ping
    pings = [P(start=6000, nframes=7, f0=1480.0, snr_db=8.0, text="CQ PJ4/K1ABC"),
             P(start=40000, nframes=20, f0=1530.0, snr_db=14.0, text="K1ABC W9XYZ EN37"),
             P(start=n // 2 + 9000, nframes=7, f0=1500.0, snr_db=7.0, text="<PJ4/K1ABC> W9XYZ RR73"),
             P(start=n - 30000, nframes=7, f0=1460.0, snr_db=6.0, text="W9XYZ <PJ4/K1ABC> RRR"),
             P(start=n - 14000, nframes=5, f0=1520.0, snr_db=3.0, text="CQ K1ABC FN42")]
    return synth.audio_stream(pings, n, seed=31)


@pytest.mark.gpu
@pytest.mark.parametrize("parts", [2, 3, 7])
def test_time_chunks_replayed_in_order_equal_the_undivided_decode(gpu, synth, parts):
    """SURVEY 8e "Partitioning": GPU stages per time piece, ONE ordered host pass.  Lines byte-identical to the
    undivided decode: snr (a recurrence over all earlier blocks) and hashed call signs (learnt in an earlier piece)
    included; LDPC successes identical too."""
    x = hashed_call_stream(synth)
    nb = (x.size - 5184) // 2592 + 1
    with gpu.Decoder(compat_cache=True, **KW) as d:
        whole, whole_res = d.decode_host(x)
        pieces = [d.decode_chunk_host(x[2592 * b0: 2592 * (b1 - 1) + 5184], b0) for b0, b1 in gpu.chunk_ranges(nb, parts)]
    texts = [l["text"] for l in whole]
    assert "CQ PJ4/K1ABC" in texts and "<PJ4/K1ABC> W9XYZ RR73" in texts and "W9XYZ <PJ4/K1ABC> RRR" in texts
    assert len({l["snr"] for l in whole}) > 2
    res = np.concatenate([p[0] for p in pieces])
    powers = np.concatenate([p[1] for p in pieces])
    assert powers.shape == (nb, 8) and res.tobytes() == whole_res.tobytes()
    assert gpu.replay_channel(powers, res, compat_cache=True) == whole
    # the second half alone does not know the call sign: order matters, and the replay is what supplies it
    late = gpu.replay_channel(pieces[-1][1], pieces[-1][0], first_block=gpu.chunk_ranges(nb, parts)[-1][0], compat_cache=True)
    assert any("<...>" in l["text"] for l in late)


@pytest.mark.gpu
@pytest.mark.parametrize("members,nchan", [(2, 1), (5, 1), (5, 2)])
def test_pool_splits_channels_in_time_when_members_outnumber_them(gpu, synth, members, nchan):
    x = hashed_call_stream(synth)
    xs = np.concatenate([x, x[::-1].copy()][:nchan])
    with gpu.Decoder(compat_cache=True, **KW) as d:
        whole, whole_res = d.decode_host(xs, nchan=nchan)
    with gpu.Pool([0] * members, compat_cache=True, **KW) as p:
        lines, res = p.decode_host(xs, nchan, result_cap=1 << 16)
        busy = [gpu.StageTimes() for _ in range(members)]
        for i, t in enumerate(busy):
            gpu.lib().msk144_get_stage_times(gpu.lib().msk144_pool_member(p._h, i), __import__("ctypes").byref(t))
    assert len(whole) > 3 and lines == whole and res.tobytes() == whole_res.tobytes()
    assert all(t.blocks > 0 for t in busy), [t.blocks for t in busy]     # every member took a share of the blocks

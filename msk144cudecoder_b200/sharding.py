"""Multi-GPU sharding of the decode path across PROCESSES.

Channels are independent units: each rank decodes its contiguous share on its own GPU and the decode lines are
merged on rank 0 in (channel, block) order.  There is no collective on the data path -- torch.distributed is
used only to gather the few bytes of text at the end, the host-side gather the north star asks for.  Works with
any backend (nccl on the GPUs, gloo in CPU tests).  Inside ONE process the same partition is done natively by
msk144_pool_* (csrc/host/pool.cpp, `msk144cudecoder_b200.Pool`).

One stream CAN also be split in time across ranks (`decode_time_sharded`): on the device a block needs only its
own 5184 samples (consecutive blocks overlap by half, main.cu:284-288).  What a stream carries from block to block
lives on the host -- the SNR noise-floor recurrence (snr_tracker.cu:39-50) and the call-sign hash tables of the text
decoder -- so the ranks run only the GPU stages of their block range and return (sub-block powers, LDPC successes);
rank 0 replays SNR and text once, in block order (`replay_channel`), and prints the sequential decoder's lines,
`snr=` values and `<...>` fields included.  Natively in one process: msk144_pool_decode_host with fewer channels
than members.
"""
from __future__ import annotations

from typing import Callable, List, Sequence


def channels_of_rank(nchan: int, rank: int, world: int) -> List[int]:
    """Contiguous, balanced partition: the first nchan % world ranks get one extra channel."""
    base, extra = divmod(nchan, world)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


def _order(lines):
    return sorted(lines, key=lambda l: (l["channel"], l["block"], l["text"]))


def gather_lines(mine: List[dict], rank: int, world: int) -> List[dict]:
    """Rank 0 gets every rank's lines merged in (channel, block) order; other ranks get their own share."""
    if world == 1:
        return _order(mine)
    import torch.distributed as dist
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(mine, bucket, dst=0)
    if rank != 0:
        return _order(mine)
    return _order([l for part in bucket for l in part])


def decode_sharded(decode_channel: Callable[[int], Sequence[dict]], nchan: int, rank: int, world: int,
                   gather: bool = True):
    """Run `decode_channel(c)` (returns decode-line dicts) for each of this rank's channels, one call per channel."""
    mine = []
    for c in channels_of_rank(nchan, rank, world):
        for l in decode_channel(c):
            mine.append(dict(l, channel=c))
    return gather_lines(mine, rank, world) if gather else _order(mine)


def decode_sharded_batch(decode_channels: Callable[[List[int]], Sequence[dict]], nchan: int, rank: int, world: int,
                         gather: bool = True):
    """Same, with ONE call for the rank's whole share: `decode_channels(channels)` decodes them as one batch (e.g.
    `Decoder.decode_host(samples_of(channels), nchan=len(channels))`) and returns lines whose `channel` is the
    index into `channels`; it is mapped back to the global channel number here."""
    chans = channels_of_rank(nchan, rank, world)
    mine = [dict(l, channel=chans[l["channel"]]) for l in decode_channels(chans)] if chans else []
    return gather_lines(mine, rank, world) if gather else _order(mine)


def gather_time_pieces(powers, results, rank: int, world: int):
    """Rank 0 gets the pieces of every rank concatenated in rank (= block) order; other ranks get None."""
    import numpy as np
    if world == 1:
        return powers, results
    import torch.distributed as dist
    bucket = [None] * world if rank == 0 else None
    dist.gather_object((powers, results), bucket, dst=0)
    if rank != 0:
        return None
    return np.concatenate([b[0] for b in bucket]), np.concatenate([b[1] for b in bucket])


def decode_time_sharded(decode_chunk: Callable[[int, int], tuple], replay: Callable, nblocks: int, rank: int, world: int):
    """ONE channel of `nblocks` blocks over `world` ranks: `decode_chunk(b0, b1)` runs the GPU stages of blocks [b0, b1)
    (e.g. `lambda b0, b1: dec.decode_chunk_host(x[2592 * b0: 2592 * (b1 - 1) + 5184], b0)`) and returns (results,
    powers); `replay(powers, results)` is the ordered host pass (`msk144cudecoder_b200.replay_channel`).  Rank 0
    returns the decode lines of the whole channel, the others an empty list."""
    import numpy as np
    from . import RESULT_DTYPE, chunk_ranges
    b0, b1 = chunk_ranges(nblocks, world)[rank]
    if b1 > b0:
        results, powers = decode_chunk(b0, b1)
    else:
        results, powers = np.zeros(0, RESULT_DTYPE), np.zeros((0, 8), np.float32)
    got = gather_time_pieces(powers, results, rank, world)
    return list(replay(*got)) if got is not None else []

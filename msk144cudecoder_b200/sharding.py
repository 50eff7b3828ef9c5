"""Multi-GPU sharding of the decode path across PROCESSES.

Channels are independent units: each rank decodes its contiguous share on its own GPU and the decode lines are
merged on rank 0 in (channel, block) order.  There is no collective on the data path -- torch.distributed is
used only to gather the few bytes of text at the end, the host-side gather the north star asks for.  Works with
any backend (nccl on the GPUs, gloo in CPU tests).  Inside ONE process the same partition is done natively by
msk144_pool_* (csrc/host/pool.cpp, `msk144cudecoder_b200.Pool`).

One stream is NOT split in time across ranks: besides the half-block input overlap, a stream carries the SNR
noise-floor recurrence (snr_tracker.cu:39-50) and the call-sign hash tables of its text decoder from block to
block, so time shards would print other `snr=` values and `<...>` fields than the sequential reference.
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

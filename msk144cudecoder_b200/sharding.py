"""Multi-GPU sharding of the decode path.

Channels (and time chunks of one stream) are independent units: each rank decodes its share on its own
GPU and the decode lines are merged on rank 0 in (channel, block) order.  There is no collective on the
data path -- torch.distributed is used only to gather the few bytes of text at the end, exactly like the
host-side gather the north star asks for.  Works with any backend (nccl on the GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from typing import Callable, List, Sequence


def channels_of_rank(nchan: int, rank: int, world: int) -> List[int]:
    """Contiguous, balanced partition: the first nchan % world ranks get one extra channel."""
    base, extra = divmod(nchan, world)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


def time_chunks(nblocks: int, world: int) -> List[range]:
    """Split one stream's blocks into contiguous chunks, one per rank.  Chunk r needs samples
    [2592*first, 2592*last + 5184): a one-block halo on the input, nothing else."""
    base, extra = divmod(nblocks, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append(range(start, start + n))
        start += n
    return out


def decode_sharded(decode_channel: Callable[[int], Sequence[dict]], nchan: int, rank: int, world: int,
                   gather: bool = True):
    """Run `decode_channel(c)` (returns decode-line dicts) for this rank's channels; rank 0 gets the merged,
    ordered list of every rank's lines (other ranks get their own)."""
    mine = []
    for c in channels_of_rank(nchan, rank, world):
        for l in decode_channel(c):
            mine.append(dict(l, channel=c))
    if world == 1 or not gather:
        return sorted(mine, key=lambda l: (l["channel"], l["block"], l["text"]))
    import torch.distributed as dist
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(mine, bucket, dst=0)
    if rank != 0:
        return sorted(mine, key=lambda l: (l["channel"], l["block"], l["text"]))
    merged = [l for part in bucket for l in part]
    return sorted(merged, key=lambda l: (l["channel"], l["block"], l["text"]))

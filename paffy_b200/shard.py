"""Multi-GPU sharding of one add_mismatches job: one process per GPU, records cut into contiguous
ranges of near-equal aligned bases, packed genome replicated, outputs concatenated in rank order. No collective is on
the data path; torch.distributed only carries the final gather of finished output bytes (and the timing barrier)."""
import numpy as np

from . import lib


def my_range(weights, rank, world):
    """-> (lo, hi): this rank's contiguous record range for the given per-record weights (aligned bases)."""
    cuts = lib.partition(np.asarray(weights, dtype=np.int64), world)
    return int(cuts[rank]), int(cuts[rank + 1])


def gather_in_order(payload: bytes, rank, world):
    """Rank 0 receives every rank's output bytes and returns them concatenated in rank (= input) order."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return payload
    objs = [None] * world if rank == 0 else None
    dist.gather_object(payload, objs, dst=0)
    return b"".join(objs) if rank == 0 else None

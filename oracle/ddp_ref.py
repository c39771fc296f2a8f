"""Oracle: host-side DDP bookkeeping — bucket assignment and DistributedSampler."""
import numpy as np


def bucket_assignment(sizes_bytes, first_cap=1024 * 1024, cap=25 * 1024 * 1024):
    """Indices per bucket for tensors visited in the given order: a bucket is
    closed as soon as it holds at least its cap; the first bucket uses the small
    cap (TORCH/nn/parallel/distributed.py:831-853, 1224-1245;
    _compute_bucket_assignment_by_size with limits [first_cap, cap])."""
    buckets, cur, size, limit = [], [], 0, first_cap
    for i, s in enumerate(sizes_bytes):
        cur.append(i)
        size += s
        if size >= limit:
            buckets.append(cur)
            cur, size, limit = [], 0, cap
    if cur:
        buckets.append(cur)
    return buckets


def distributed_sampler_indices(length, world, rank, perm):
    """TORCH/utils/data/distributed.py:105-134 with drop_last=False: `perm` is the
    randperm(length) of this epoch (or arange when shuffle=False); pad by wrapping
    to a multiple of `world`, then take every world-th index from `rank`."""
    idx = list(perm)
    total = -(-length // world) * world
    pad = total - len(idx)
    if pad > 0:
        idx += (idx * -(-pad // len(idx)))[:pad]
    return np.asarray(idx[rank:total:world], dtype=np.int64)

"""Frame-parallel sharding of a batch over the GPUs of one box.

Every frame of the detector (and every minibatch image of the layer) is independent, so a batch is
split contiguously over ranks, weights are replicated, and NO data-path collective exists.  The only
exchange is the host-side gather of the variable-length detection lists, done with
`torch.distributed.all_gather_object` on whatever backend the process group has (NCCL ranks use the
group's gloo/cpu side; the tests run it on gloo).
"""


def shard_range(n_items, rank, world):
    """Contiguous [lo, hi) of `n_items` for `rank` of `world`; sizes differ by at most one, earlier ranks larger."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError('bad rank/world: %r/%r' % (rank, world))
    if n_items < 0:
        raise ValueError('n_items must be >= 0')
    base, rem = divmod(n_items, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_frames(frames, rank, world):
    """The slice of a (B, ...) batch this rank processes."""
    lo, hi = shard_range(len(frames), rank, world)
    return frames[lo:hi]


def gather_detections(local_results, group=None):
    """All ranks' per-frame results concatenated in frame order (same list on every rank).
    local_results: list (one entry per local frame) of picklable per-frame results."""
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return list(local_results)
    parts = [None] * dist.get_world_size(group)
    dist.all_gather_object(parts, list(local_results), group=group)
    out = []
    for p in parts:
        out.extend(p)
    return out


def predict_sharded(detector, images, rank, world, group=None):
    """Frame-parallel `predict_batch`: each rank runs its contiguous slice on its own GPU, results gathered on host."""
    mine = shard_frames(images, rank, world)
    res = detector.predict_batch(mine) if len(mine) else []
    return gather_detections(res, group)

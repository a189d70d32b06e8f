"""Batch sharding across the GPUs of one box: every sample is independent, so rank r owns a
contiguous slice of the batch and the forward path needs no collective.  Fitting adds one all-reduce of a few
loss scalars per step (NCCL over NVLink on GPUs; gloo in the CPU tests of this host logic)."""
import torch


def shard_range(n, rank, world):
    """Contiguous, balanced [start, stop) slice of n samples for `rank` of `world` (sizes differ by <= 1)."""
    if not (0 <= rank < world):
        raise ValueError('rank %d outside world %d' % (rank, world))
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def shard_sizes(n, world):
    return [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]


def allreduce_scalars(t, group=None):
    """Sum a small tensor of loss scalars over ranks in place (no-op without an initialised process group).
    Each rank's loss share is already divided by the GLOBAL element count (smplb_fit_io.n_total), so the sum is
    the reference's mean over the whole batch (pose_model_3d.py:233-235)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def global_loss(fitter_loss, group=None):
    """[uncapped share, capped, gate] of one rank -> global uncapped loss (device tensor, 1 element)."""
    return allreduce_scalars(fitter_loss[:1].clone(), group)

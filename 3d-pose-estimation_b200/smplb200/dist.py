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


class SideStreamAllReduce:
    """Loss-scalar all-reduce off the critical path: the sum of step k runs on a side stream while
    the kernels of step k+1 run on the main stream; the value is only consumed for logging / stop criteria.  Two
    slots alternate, so the main stream waits for the reduction of step k-2 (long finished) before it overwrites
    that slot."""

    def __init__(self, n, device, group=None):
        self.buf = torch.zeros((2, n), dtype=torch.float32, device=device)
        self.group = group
        self.k = 0
        self.cuda = torch.device(device).type == 'cuda'
        self.side = torch.cuda.Stream(device=device) if self.cuda else None
        self.done = [None, None]

    def push(self, scalars):
        """Queue the all-reduce of `scalars` (n floats on the device); returns the slot tensor it will land in."""
        slot = self.k & 1
        self.k += 1
        if not self.cuda:
            self.buf[slot].copy_(scalars)
            allreduce_scalars(self.buf[slot], self.group)
            return self.buf[slot]
        main = torch.cuda.current_stream(self.buf.device)
        if self.done[slot] is not None:
            main.wait_event(self.done[slot])
        self.buf[slot].copy_(scalars)
        self.side.wait_stream(main)
        with torch.cuda.stream(self.side):
            allreduce_scalars(self.buf[slot], self.group)
            ev = torch.cuda.Event()
            ev.record(self.side)
        self.done[slot] = ev
        return self.buf[slot]

    def latest(self):
        """Synchronise with the most recent reduction and return its value (host sync only here)."""
        slot = (self.k - 1) & 1
        if self.cuda and self.done[slot] is not None:
            self.done[slot].synchronize()
        return self.buf[slot]

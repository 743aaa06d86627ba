"""Host-side sharding of the hot path over the GPUs of one node.

Gain calculation, the convolver bank and HOA shard by objects / channels with no data-path
collective; object-sharded mixing has the one exchange step: every rank renders its objects into a
partial [M x T] loudspeaker bus and the buses are summed (`reduce_bus`, one all-reduce; NCCL over
NVLink on the GPUs, gloo in the CPU tests).  One process per GPU, `torch.distributed` plumbing."""


def shard_range(n_units, rank, world):
    """Contiguous, balanced [start, end) of `n_units` objects / channels for `rank`: the first
    n_units % world ranks hold one unit more."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank outside the world")
    base, extra = divmod(int(n_units), world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def object_groups(n_objects, max_group=256):
    """Groups of <= max_group objects a rank mixes one after the other into its bus (the render
    kernel's tensor-core path takes up to 256 input channels per launch)."""
    return [(g, min(n_objects, g + max_group)) for g in range(0, n_objects, max_group)]


def reduce_bus(bus, group=None):
    """Sum the per-rank partial loudspeaker buses in place on every rank (no-op for one rank)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(bus, op=dist.ReduceOp.SUM, group=group)
    return bus

"""Host-side sharding of the hot path over the GPUs of one node.

Gain calculation, the convolver bank and HOA shard by objects / channels with no data-path
collective; object-sharded mixing has the one exchange step: every rank renders its objects into a
partial [M x T] loudspeaker bus and the buses are summed.  On the GPUs that step is part of the
product: the node-wide bus of the C ABI (libear_b200.Bus: peer stores from the render epilogue, a
deterministic sum per time slice; `connect_bus` hands the IPC handles around).  `reduce_bus` /
`mix_pipelined` are the torch.distributed all-reduce form (gloo in the CPU tests, NCCL as the
comparison arm of bench.py).  One process per GPU, `torch.distributed` plumbing."""


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


def bus_slices(n_samples, world, tile=256):
    """The time slices [t0, t1) of a node-wide bus (libear_b200.Bus / earb200_bus_slice): rank r owns
    slice r; slices are whole render tiles, the last ones may be short or empty."""
    per = -(-int(n_samples) // world)
    per = -(-per // tile) * tile
    return [(min(n_samples, r * per), min(n_samples, (r + 1) * per)) for r in range(world)]


def connect_bus(bus, group=None):
    """One process per GPU: hand the ranks' IPC handles around (any transport would do: here the
    process group's all_gather_object) and open them.  The exchange itself never touches
    torch.distributed or NCCL: it is stores over NVLink from the render kernel's epilogue."""
    import torch.distributed as dist
    if bus.n_ranks == 1:
        return bus
    handles = [None] * bus.n_ranks
    dist.all_gather_object(handles, bus.ipc_handle(), group=group)
    bus.open_ipc(handles)
    dist.barrier(group=group)
    return bus


def reduce_bus(bus, group=None):
    """Sum the per-rank partial loudspeaker buses in place on every rank (no-op for one rank)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(bus, op=dist.ReduceOp.SUM, group=group)
    return bus


def chunk_ranges(n_samples, n_chunks, align=512):
    """[start, end) sample ranges of at most n_chunks chunks whose boundaries are multiples of
    `align` (whole render tiles / convolver blocks); the last chunk takes the remainder."""
    n_samples, n_chunks = int(n_samples), max(1, int(n_chunks))
    per = -(-n_samples // n_chunks)
    per = max(align, -(-per // align) * align)
    return [(t, min(n_samples, t + per)) for t in range(0, n_samples, per)]


def reduce_bus_rows(bus, t0, t1, group=None):
    """All-reduce the columns [t0, t1) of a planar [M x T] bus in place: the M row slices are
    contiguous, so they go out as one coalesced NCCL group (one fused launch)."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1):
        return
    rows = [bus[m, t0:t1] for m in range(bus.shape[0])]
    if bus.is_cuda and hasattr(dist, "_coalescing_manager"):
        with dist._coalescing_manager(group=group, device=bus.device):
            for r in rows:
                dist.all_reduce(r, op=dist.ReduceOp.SUM, group=group)
    else:
        for r in rows:
            dist.all_reduce(r, op=dist.ReduceOp.SUM, group=group)


def chunked_bus(n_out, n_samples, n_chunks, device=None, dtype=None, align=512):
    """A loudspeaker bus stored chunk-major, [n_chunks][n_out][chunk_samples]: every time chunk is
    one contiguous block, so its all-reduce is a single collective.  Returns (bus, ranges); chunk k
    holds the samples ranges[k] = [t0, t1) in bus[k, :, :t1 - t0]."""
    import torch
    ranges = chunk_ranges(n_samples, n_chunks, align)
    per = ranges[0][1] - ranges[0][0]
    bus = torch.zeros((len(ranges), n_out, per), device=device, dtype=dtype or torch.float32)
    return bus, ranges


def mix_pipelined(bus, render_chunk, n_chunks=8, group=None, reduce_stream=None, align=512, ranges=None):
    """Object-sharded mixing with the exchange step hidden behind the rendering: the bus is cut
    into time chunks; chunk k is rendered on the current stream and summed over the ranks on
    `reduce_stream` while chunk k + 1 renders.  Two bus layouts:
      * chunk-major (3-D, from chunked_bus; pass its `ranges`): render_chunk(k, t0, t1) writes
        bus[k, :, :t1 - t0]; one contiguous all-reduce per chunk -- the fast form;
      * planar [M x T] (2-D): render_chunk(t0, t1) writes bus[:, t0:t1]; a chunk is M row slices,
        reduced as one coalesced group of M collectives (measured latency-bound: 112 small
        all-reduces per 60 s bus cost more than the single unchunked one).
    On one rank (or without a process group) it only renders.  Returns with both streams joined."""
    import torch
    import torch.distributed as dist
    sharded = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    chunk_major = bus.dim() == 3
    if chunk_major:
        if ranges is None or len(ranges) != bus.shape[0]:
            raise ValueError("a chunk-major bus needs the ranges chunked_bus returned")
        chunks = list(ranges)
    else:
        chunks = chunk_ranges(bus.shape[1], n_chunks if sharded else 1, align)

    def render(k):
        t0, t1 = chunks[k]
        if chunk_major:
            render_chunk(k, t0, t1)
        else:
            render_chunk(t0, t1)

    def reduce(k):
        if chunk_major:
            dist.all_reduce(bus[k], op=dist.ReduceOp.SUM, group=group)
        else:
            reduce_bus_rows(bus, chunks[k][0], chunks[k][1], group)

    if not (sharded and bus.is_cuda):
        for k in range(len(chunks)):
            render(k)
            if sharded:
                reduce(k)
        return bus
    main = torch.cuda.current_stream(bus.device)
    side = reduce_stream or torch.cuda.Stream(bus.device)
    for k in range(len(chunks)):
        render(k)
        done = torch.cuda.Event()
        done.record(main)
        side.wait_event(done)
        with torch.cuda.stream(side):
            reduce(k)
    main.wait_stream(side)
    return bus

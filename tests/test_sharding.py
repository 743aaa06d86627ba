"""Multi-GPU plan on CPU: world-size-2 gloo processes.  Objects shard contiguously with
no collective; object-sharded mixing sums per-rank partial buses with one all-reduce.  The per-rank
rendering is done by the numpy oracle here (test infrastructure) -- what is under test is the
sharding arithmetic and the reduce, the same code bench.py drives over NCCL."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from libear_b200.sharding import (bus_slices, chunk_ranges, chunked_bus, mix_pipelined, object_groups, reduce_bus,
                                   shard_range)


def test_shard_range_is_a_partition():
    for n in (0, 1, 7, 512, 4096, 1_000_003):
        for world in (1, 2, 3, 8):
            parts = [shard_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_object_groups():
    assert object_groups(512) == [(0, 256), (256, 512)]
    assert object_groups(300) == [(0, 256), (256, 300)]
    assert object_groups(0) == []


def test_chunk_ranges():
    assert chunk_ranges(2_880_000, 8) == [(t, min(2_880_000, t + 360_448)) for t in range(0, 2_880_000, 360_448)]
    assert chunk_ranges(1000, 4) == [(0, 512), (512, 1000)]
    assert chunk_ranges(100, 8) == [(0, 100)]
    for T, k in ((4096 + 37, 3), (512 * 7, 7), (1, 1)):
        c = chunk_ranges(T, k)
        assert c[0][0] == 0 and c[-1][1] == T and len(c) <= k
        assert all(a[1] == b[0] for a, b in zip(c, c[1:])) and all(a % 512 == 0 for a, _ in c)


def test_bus_slices_partition_the_bus_in_whole_tiles():
    assert bus_slices(2_880_000, 8) == [(r * 360_192, min(2_880_000, (r + 1) * 360_192)) for r in range(8)]
    assert bus_slices(4096, 2) == [(0, 2048), (2048, 4096)]
    assert bus_slices(300, 4) == [(0, 256), (256, 300), (300, 300), (300, 300)]
    for T, w in ((9572, 2), (1, 8), (2_880_000, 4), (256 * 37 + 100, 4)):
        sl = bus_slices(T, w)
        assert sl[0][0] == 0 and sl[-1][1] == T and all(a[1] == b[0] for a, b in zip(sl, sl[1:]))
        assert all(a % 256 == 0 for a, _ in sl if a < T)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import ear_oracle as eo
        rng = np.random.default_rng(7)          # same seed on every rank: the global problem
        N, M, T = 10, 6, 1500
        x = rng.normal(0, 0.25, (N, T)).astype(np.float32)
        samples = np.array([0, 400, 400, 1100, 2000])
        gains = rng.uniform(0, 1, (len(samples), N, M)).astype(np.float32)
        a, b = shard_range(N, rank, world)
        partial = np.zeros((M, T), dtype=np.float32)
        for g0, g1 in object_groups(b - a, max_group=3):          # groups accumulate into the bus
            partial += eo.interp_matrix(x[a + g0:a + g1], samples, gains[:, a + g0:a + g1])
        bus = torch.from_numpy(partial.copy())
        reduce_bus(bus)
        full = eo.interp_matrix(x, samples, gains)
        err = float(np.abs(bus.numpy() - full).max() / np.abs(full).max())
        # the same mix chunk by chunk, each chunk summed over the ranks as soon as it is rendered
        bus2 = torch.zeros((M, T), dtype=torch.float32)

        def render_chunk(t0, t1):
            bus2[:, t0:t1] = torch.from_numpy(eo.interp_matrix(x[a:b, t0:t1], samples, gains[:, a:b], block_start=t0))
        mix_pipelined(bus2, render_chunk, n_chunks=3)
        err = max(err, float(np.abs(bus2.numpy() - full).max() / np.abs(full).max()))
        # and with the chunk-major bus (one contiguous all-reduce per chunk)
        bus3, ranges = chunked_bus(M, T, 3)

        def render_chunk3(k, t0, t1):
            bus3[k, :, :t1 - t0] = torch.from_numpy(
                eo.interp_matrix(x[a:b, t0:t1], samples, gains[:, a:b], block_start=t0))
        mix_pipelined(bus3, render_chunk3, ranges=ranges)
        planar = np.concatenate([bus3[k, :, :t1 - t0].numpy() for k, (t0, t1) in enumerate(ranges)], axis=1)
        err = max(err, float(np.abs(planar - full).max() / np.abs(full).max()))
        # gains shard with no collective: each rank's rows are just a slice of the global answer
        pan = eo.PointSourcePanner("0+5+0")
        dirs = rng.normal(size=(64, 3))
        sl = slice(*shard_range(64, rank, world))
        mine, mine_region, _ = pan.handle(dirs[sl])
        full_gains, full_region, _ = pan.handle(dirs)
        ok_slice = bool(np.array_equal(mine, full_gains[sl]) and np.array_equal(mine_region, full_region[sl]))
        # the IPC handles of a node-wide bus travel through the process group (connect_bus): every
        # rank ends with all handles in rank order
        from libear_b200.sharding import connect_bus

        class FakeBus:
            n_ranks = world

            def ipc_handle(self):
                return bytes([rank]) * 64

            def open_ipc(self, handles):
                self.opened = handles
        fb = connect_bus(FakeBus())
        ok_slice = ok_slice and fb.opened == [bytes([r]) * 64 for r in range(world)]
        np.save(os.path.join(out_dir, f"r{rank}.npy"), np.array([err, float(ok_slice), a, b]))
    finally:
        dist.destroy_process_group()


def test_object_sharded_mixing_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    res = [np.load(tmp_path / f"r{r}.npy") for r in range(world)]
    assert [tuple(r[2:]) for r in res] == [(0.0, 5.0), (5.0, 10.0)]
    for r in res:
        assert r[0] < 1e-6      # summed partial buses == the full render (within the audio tolerance)
        assert r[1] == 1.0

"""The exchange step of object-sharded mixing ON GPUs: objects sharded over the ranks of
a node-wide bus (earb200_bus_*), every rank's render storing its tiles into the owners' slots from
the kernel's epilogue, the owners summing -- against ONE interpolator rendering all objects on one
GPU (libear's contract: "summed with other objects into an n-channel bus",
ref include/ear/gain_calculators.hpp:43-56) and against libear itself.

The ranks live in this process (earb200_bus_connect_local); with a single GPU on the box they all
sit on device 0 -- same code path (the slots are then local instead of peer memory) -- and with
two or more GPUs the same tests also run one rank per device over NVLink.  The multi-process form
(IPC handles) is exercised by bench.py under torchrun.  Tolerance: 1e-6 of peak (the sum over the
ranks is taken in another order than the single interpolator's)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def device_sets():
    import torch
    sets = [("one_device", [0, 0]), ("one_device_x4", [0, 0, 0, 0])]
    n = torch.cuda.device_count()
    if n >= 2:
        sets.append(("two_devices", [0, 1]))
    if n >= 4:
        sets.append(("four_devices", [0, 1, 2, 3]))
    return sets


def make_case(N, M, T, seed, spacing=512):
    rng = np.random.default_rng(seed)
    x = rng.normal(0, 0.25, (N, T)).astype(np.float32)
    samples = np.arange(0, T + spacing, spacing, dtype=np.int64)
    gains = (rng.uniform(0, 1, (len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
    return x, samples, gains


@pytest.mark.parametrize("name,devices", device_sets())
@pytest.mark.parametrize("mode", ["fused", "accumulate_groups", "ragged_push", "unaligned_points",
                                  "unaligned_points_accumulate"])
def test_sharded_mix_equals_single_gpu_render(eb, ref, name, devices, mode):
    import torch
    R = len(devices)
    N, M = 64 * R, 12
    T = 256 * 37 if mode != "ragged_push" else 256 * 37 + 100
    # unaligned_points: interpolation points every 1,000 samples, so segments end inside render tiles:
    # the tiles go to the bus from the kernel's epilogue, the ragged runs between them are pushed
    x, samples, gains = make_case(N, M, T, 40 + R, spacing=1000 if mode.startswith("unaligned") else 512)
    whole = eb.GainInterpolator(N, M, device=0)
    whole.set_points(samples, gains)
    exp = whole.process(0, x)
    peak = np.abs(exp).max()
    assert np.abs(exp - ref.interp_matrix(x, samples, gains)).max() / peak < 2e-6   # N up to 256 float terms

    buses = [eb.Bus(M, T, rank=r, n_ranks=R, device=d) for r, d in enumerate(devices)]
    eb.Bus.connect_local(buses)
    per = N // R
    for epoch in range(3):   # the slot sets alternate: later epochs must not see earlier data
        scale = np.float32(1.0 + epoch)
        keep, outs = [], []
        for r, d in enumerate(devices):
            dev = torch.device("cuda", d)
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream(dev).cuda_stream
                xs = torch.from_numpy(x[r * per:(r + 1) * per] * scale).to(dev)
                keep.append(xs)
                if mode in ("accumulate_groups", "unaligned_points_accumulate"):
                    half = per // 2
                    g0 = eb.GainInterpolator(half, M, device=d)
                    g0.set_points(samples, gains[:, r * per:r * per + half])
                    g1 = eb.GainInterpolator(per - half, M, device=d)
                    g1.set_points(samples, gains[:, r * per + half:(r + 1) * per])
                    g1.set_accumulate(True)
                    part = torch.full((M, T), float("nan"), dtype=torch.float32, device=dev)
                    g0.process_device(0, T, xs.data_ptr(), T, part.data_ptr(), T, stream=stream)
                    g1.process_device_bus(0, T, xs[half].data_ptr(), T, buses[r], 0, d_partial=part.data_ptr(),
                                          partial_stride=T, stream=stream)
                    keep += [g0, g1, part]
                else:
                    gi = eb.GainInterpolator(per, M, device=d)
                    gi.set_points(samples, gains[:, r * per:(r + 1) * per])
                    part = None
                    if mode == "unaligned_points":
                        part = torch.full((M, T), float("nan"), dtype=torch.float32, device=dev)
                    if mode == "ragged_push":
                        part = torch.full((M, T + 4), float("nan"), dtype=torch.float32, device=dev)
                    gi.process_device_bus(0, T, xs.data_ptr(), T, buses[r], 0,
                                          d_partial=None if part is None else part.data_ptr(),
                                          partial_stride=0 if part is None else part.shape[1], stream=stream)
                    keep += [gi, part]
        # several ranks share a device (and its stream) here: all publish before any waits
        for r, d in enumerate(devices):
            with torch.cuda.device(d):
                buses[r].publish(stream=torch.cuda.current_stream(torch.device("cuda", d)).cuda_stream)
        for r, d in enumerate(devices):
            dev = torch.device("cuda", d)
            with torch.cuda.device(dev):
                t0, t1 = buses[r].slice()
                sl = torch.full((M, max(1, t1 - t0)), float("nan"), dtype=torch.float32, device=dev)
                outs.append((t0, t1, sl))
                buses[r].reduce(d_out=sl.data_ptr(), out_stride=sl.shape[1], gather=True,
                                stream=torch.cuda.current_stream(dev).cuda_stream)
        for r, d in enumerate(devices):
            with torch.cuda.device(d):
                buses[r].wait_result(stream=torch.cuda.current_stream(torch.device("cuda", d)).cuda_stream)
        for d in set(devices):
            torch.cuda.synchronize(d)
        want = exp * scale
        for r, d in enumerate(devices):
            t0, t1, sl = outs[r]
            got = sl.cpu().numpy()[:, :t1 - t0]
            assert np.abs(got - want[:, t0:t1]).max() / (peak * scale) < 1e-6, (name, mode, epoch, r, "slice")
            # the gathered full bus of every rank
            full = torch.full((M, T), float("nan"), dtype=torch.float32, device=torch.device("cuda", d))
            with torch.cuda.device(d):
                buses[r].copy_result(full.data_ptr(), T, 0, T, stream=torch.cuda.current_stream(full.device).cuda_stream)
                torch.cuda.synchronize(d)
            assert np.abs(full.cpu().numpy()[:, :T] - want).max() / (peak * scale) < 1e-6, (name, mode, epoch, r, "full")


def test_bus_argument_checks(eb):
    with pytest.raises(eb.InvalidArgument):
        eb.Bus(12, 1000, rank=2, n_ranks=2)
    with pytest.raises(eb.InvalidArgument):
        eb.Bus(12, 1000, rank=0, n_ranks=9)
    b = eb.Bus(12, 4096, rank=0, n_ranks=2)
    with pytest.raises(eb.InvalidArgument):      # not connected yet
        b.reduce(gather=True)
    assert b.slice(0) == (0, 2048) and b.slice(1) == (2048, 4096)
    assert len(b.ipc_handle()) == 64
    gi = eb.GainInterpolator(4, 5)
    one = eb.Bus(12, 4096)
    with pytest.raises(eb.InvalidArgument):      # channel count of the bus
        gi.process_device_bus(0, 256, 1024, 256, one, 0)

"""earb200_objects_calculate_device_static: the objects gain calculator with its scratch in a
file-scope device array (no cudaMalloc in the call), checked through the C ABI against
earb200_objects_calculate_device on the same inputs.

The entry point has no counterpart in libear.  This
file is named to be collected last.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-5  # the tolerance the objects parity tests use (float extent core); same kernels, so ~0 here


def _metadata(rng, n):
    az = rng.uniform(-180, 180, n)
    el = np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
    dist = rng.uniform(0.2, 1.0, n)
    # a mix of point sources (no extent), small extents (both branches) and wide ones
    kind = rng.integers(0, 4, n)
    w = np.where(kind == 0, 0.0, np.where(kind == 1, rng.uniform(0, 8, n), rng.uniform(0, 360, n)))
    h = np.where(kind == 0, 0.0, np.where(kind == 1, rng.uniform(0, 8, n), rng.uniform(0, 180, n)))
    d = np.where(rng.random(n) < 0.4, rng.uniform(0, 1, n), 0.0)
    gain = rng.uniform(0.1, 1.0, n)
    diffuse = rng.uniform(0, 1, n)
    return np.stack([az, el, dist, w, h, d, gain, diffuse])


def _both(eb, gc, cols, out_is_double, with_regions):
    """(existing entry point, static entry point) -> ((direct, diffuse, regions) x 2)"""
    import torch
    dev = torch.device("cuda", 0)
    n = cols.shape[1]
    M = gc.num_outputs
    dt = torch.float64 if out_is_double else torch.float32
    d_in = torch.from_numpy(cols).to(dev)
    ptrs = [d_in[j].data_ptr() for j in range(8)]
    stream = torch.cuda.current_stream().cuda_stream
    outs = []
    for fn in (gc.calculate_device, gc.calculate_device_static):
        # poisoned outputs: every element must be written by the call
        direct = torch.full((n, M), float("nan"), dtype=dt, device=dev)
        diffuse = torch.full((n, M), float("nan"), dtype=dt, device=dev)
        regions = torch.full((n,), -7, dtype=torch.int32, device=dev) if with_regions else None
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(0)[0]
        fn(n, ptrs, direct.data_ptr(), diffuse.data_ptr(), regions.data_ptr() if with_regions else None,
           out_is_double=out_is_double, stream=stream)
        torch.cuda.synchronize()
        free_after = torch.cuda.mem_get_info(0)[0]
        outs.append((direct.cpu().numpy(), diffuse.cpu().numpy(),
                     regions.cpu().numpy() if with_regions else None, free_before - free_after))
    return outs


def _check(old, new):
    assert np.isfinite(new[0]).all() and np.isfinite(new[1]).all()
    assert np.abs(new[0] - old[0]).max() <= TOL
    assert np.abs(new[1] - old[1]).max() <= TOL
    if old[2] is not None:
        assert np.array_equal(new[2], old[2])


def test_static_scratch_entry_point_matches_the_allocating_one_on_two_calls_in_a_row(eb):
    rng = np.random.default_rng(20261105)
    gc = eb.GainCalculatorObjects("9+10+3")
    # call 1 fills the whole array (the maximum batch, every object with a region index) ...
    n1 = gc.STATIC_MAX
    old1, new1 = _both(eb, gc, _metadata(rng, n1), out_is_double=False, with_regions=True)
    _check(old1, new1)
    # ... call 2, right after it and smaller, finds call 1's contents in the array
    old2, new2 = _both(eb, gc, _metadata(rng, 777), out_is_double=True, with_regions=False)
    _check(old2, new2)
    # the kernels of call 1 are loaded by now: the second static call took no device memory
    assert new2[3] == 0, f"device memory changed by {new2[3]} bytes across a call that must not allocate"
    # a third, on another layout / handle, shares the same array
    gc2 = eb.GainCalculatorObjects("0+5+0")
    old3, new3 = _both(eb, gc2, _metadata(rng, 2049), out_is_double=False, with_regions=True)
    _check(old3, new3)
    assert new3[3] == 0


def test_static_scratch_entry_point_rejects_batches_beyond_its_maximum(eb):
    import torch
    gc = eb.GainCalculatorObjects("4+5+0")
    n = gc.STATIC_MAX + 1
    dev = torch.device("cuda", 0)
    d_in = torch.zeros((8, n), dtype=torch.float64, device=dev)
    out = torch.zeros((2, n, gc.num_outputs), dtype=torch.float32, device=dev)
    ptrs = [d_in[j].data_ptr() for j in range(8)]
    with pytest.raises(Exception) as e:
        gc.calculate_device_static(n, ptrs, out[0].data_ptr(), out[1].data_ptr())
    assert "EARB200_OBJECTS_STATIC_MAX" in str(e.value) and str(gc.STATIC_MAX) in str(e.value)
    # an empty batch is fine, as with the allocating entry point
    gc.calculate_device_static(0, ptrs, out[0].data_ptr(), out[1].data_ptr())

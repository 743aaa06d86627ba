"""K3 parity: GainInterpolator vs the compiled libear reference.

Mirrors tests/gain_interpolator_tests.cpp (basic/step/only_step/one_point :104-183,
vector :187-219, matrix :221-257).  Tolerance: rendered audio within 1e-6 of peak
(north star; the kernel re-associates the per-sample gain ramp, see render_kernels.cu).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def peak_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def run_both(eb, ref, x, samples, gains, block_start=0, call_size=None):
    N, T = x.shape
    M = gains.shape[2]
    exp = ref.interp_matrix(x, samples, gains, block_start0=block_start, call_size=call_size)
    gi = eb.GainInterpolator(N, M)
    gi.set_points(samples, gains)
    cs = call_size or T
    out = np.concatenate([gi.process(block_start + t, x[:, t:t + cs]) for t in range(0, T, cs)], axis=1)
    return out, exp


@pytest.mark.parametrize("call_size", [None, 50, 75, 100, 512, 1000])
def test_matrix_segments(eb, ref, call_size):
    rng = np.random.default_rng(0)
    N, M, T = 5, 3, 2000
    x = rng.normal(size=(N, T)).astype(np.float32)
    samples = [100, 400, 400, 900, 1300, 1300, 1301, 1700]  # ramps, steps, a 1-sample ramp
    gains = rng.uniform(-1, 1, size=(len(samples), N, M)).astype(np.float32)
    gains[4] = gains[3]  # constant segment between equal points
    out, exp = run_both(eb, ref, x, samples, gains, call_size=call_size)
    assert peak_err(out, exp) < 1e-6


def test_block_start_offsets_and_negative_indices(eb, ref):
    rng = np.random.default_rng(1)
    N, M, T = 3, 4, 777
    x = rng.normal(size=(N, T)).astype(np.float32)
    samples = [-500, -100, 300, 1200]
    gains = rng.uniform(0, 1, size=(4, N, M)).astype(np.float32)
    for start in [-600, -100, 0, 13, 1000, 5000]:
        out, exp = run_both(eb, ref, x, samples, gains, block_start=start, call_size=128)
        assert peak_err(out, exp) < 1e-6, start


def test_one_point_and_only_step(eb, ref):
    rng = np.random.default_rng(2)
    x = rng.normal(size=(2, 300)).astype(np.float32)
    g = rng.uniform(0, 1, size=(1, 2, 2)).astype(np.float32)
    out, exp = run_both(eb, ref, x, [100], g)
    assert np.array_equal(out, exp) or peak_err(out, exp) < 1e-6
    g2 = rng.uniform(0, 1, size=(2, 2, 2)).astype(np.float32)
    out, exp = run_both(eb, ref, x, [150, 150], g2, call_size=100)
    assert peak_err(out, exp) < 1e-6


def exact_render(x, samples, gains):
    """float64 evaluation of LinearInterpMatrix (ref include/ear/dsp/gain_interpolator.hpp:248-300) for
    strictly increasing sample indices that cover the whole call."""
    x, gains = x.astype(np.float64), gains.astype(np.float64)
    T = x.shape[1]
    out = np.zeros((gains.shape[2], T))
    for i in range(len(samples) - 1):
        a, b = int(samples[i]), int(samples[i + 1])
        t = np.arange(max(a, 0), min(b, T))
        if len(t) == 0:
            continue
        p = (t - a) / float(b - a)
        out[:, t] = (np.einsum("nm,nt->mt", gains[i], x[:, t]) * (1.0 - p) +
                     np.einsum("nm,nt->mt", gains[i + 1], x[:, t]) * p)
    return out


@pytest.mark.parametrize("N,M", [(1, 1), (1, 7), (16, 12), (32, 24), (7, 9), (3, 17), (2, 32), (128, 14),
                                 (128, 24), (20, 31), (256, 22), (300, 12), (520, 24)])
def test_channel_shapes(eb, ref, N, M):
    # (1,1) = LinearInterpSingle, (1,M) = LinearInterpVector, (16,12) = HOA order 3 -> 4+7+0,
    # (32,24) = the reference's matrix_benchmark shape, (128,14) = BASELINE config 2 shape;
    # M > 16 runs as two groups of output channels per tile on the tensor-core path, N > 256 as
    # groups of 256 inputs accumulated into the output
    rng = np.random.default_rng(N * 100 + M)
    T = 4096 + 37
    x = np.clip(rng.normal(0, 0.25, size=(N, T)), -1, 1).astype(np.float32)
    samples = np.arange(0, T + 512, 512)
    gains = (rng.uniform(0, 1, size=(len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
    out, exp = run_both(eb, ref, x, samples, gains)
    if N < 128:
        assert peak_err(out, exp) < 1e-6
    else:
        # 128+ products per sample: libear's sequential fp32 sum and the tensor-core fp32 sum each sit
        # ~5e-7 of peak away from the exact result, so the bar is 1e-6 against the float64 evaluation
        # of the same formula and 2e-6 between the two fp32 paths
        exact = exact_render(x, samples, gains)
        assert peak_err(out, exact) < 1e-6
        assert peak_err(exp, exact) < 1.5e-6  # the reference's own fp32 rounding
        assert peak_err(out, exp) < 2e-6
    if N == 1:
        v = ref.interp_vector(x[0], samples, gains[:, 0, :])
        assert peak_err(out, v) < 1e-6
    if N == 1 and M == 1:
        s = ref.interp_single(x[0], samples, gains[:, 0, 0])
        assert peak_err(out[0], s) < 1e-6


def test_unaligned_buffers(eb, ref):
    # process() on sub-ranges that are not 16-byte aligned relative to the segment grid
    rng = np.random.default_rng(3)
    N, M, T = 4, 6, 1500
    x = rng.normal(size=(N, T)).astype(np.float32)
    samples = [3, 258, 1029]
    gains = rng.uniform(0, 1, size=(3, N, M)).astype(np.float32)
    for cs in [1, 3, 127, 333]:
        out, exp = run_both(eb, ref, x[:, :400], samples, gains, call_size=cs)
        assert peak_err(out, exp) < 1e-6, cs


def test_errors(eb):
    gi = eb.GainInterpolator(2, 2)
    x = np.zeros((2, 10), dtype=np.float32)
    with pytest.raises(eb.InvalidArgument):
        gi.process(0, x)  # no points
    gi.set_points([10, 5], np.zeros((2, 2, 2), dtype=np.float32))
    with pytest.raises(eb.InvalidArgument, match="not sorted"):
        gi.process(0, x)


def test_linearity_full_size_channels(eb):
    # size-independent property at the BASELINE shape (128 -> 14, 512-sample ramps):
    # render(a*x + y) == a*render(x) + render(y)
    rng = np.random.default_rng(4)
    N, M, T = 128, 14, 512 * 40
    x = rng.normal(0, 0.25, size=(N, T)).astype(np.float32)
    y = rng.normal(0, 0.25, size=(N, T)).astype(np.float32)
    samples = np.arange(0, T + 512, 512)
    gains = (rng.uniform(0, 1, size=(len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
    gi = eb.GainInterpolator(N, M)
    gi.set_points(samples, gains)
    rx, ry, rxy = gi.process(0, x), gi.process(0, y), gi.process(0, 0.5 * x + y)
    assert np.abs(rxy - (0.5 * rx + ry)).max() / np.abs(rxy).max() < 1e-5


@pytest.mark.parametrize("M", [12, 24])
@pytest.mark.parametrize("T,call", [(4096 + 37, None), (3000, 700)])
def test_accumulate_mixes_channel_groups(eb, ref, T, call, M):
    """Extension earb200_interp_set_accumulate: two groups of objects rendered one after the other
    into the same bus equal one interpolator over all of them (the reference on the full matrix)."""
    rng = np.random.default_rng(21)
    N = 48
    x = rng.normal(0, 0.25, (N, T)).astype(np.float32)
    samples = np.arange(-100, T + 512, 512)
    g = (rng.uniform(0, 1, (len(samples), N, M)) / 4).astype(np.float32)
    exp = ref.interp_matrix(x, samples, g)
    a, b = eb.GainInterpolator(32, M), eb.GainInterpolator(16, M)
    a.set_points(samples, g[:, :32])
    b.set_points(samples, g[:, 32:])
    b.set_accumulate(True)
    outs = []
    step = call or T
    for t in range(0, T, step):
        bus = a.process(t, x[:32, t:t + step])
        outs.append(b.process(t, x[32:, t:t + step], out=bus))
    out = np.concatenate(outs, axis=1)
    assert np.abs(out - exp).max() / np.abs(exp).max() < 1e-6

"""OPT-IN objectDivergence / channelLock (earb200_objects_enable_extensions).  PARITY UNPINNED: libear
0.9 throws ear::not_implemented for both (ref src/object_based/gain_calculator_objects.cpp:37-43,
tests/gain_calculator_objects_tests.cpp:134-169), so there is no reference output.  What is checked:
  * the default stays libear's: the features fail with NotImplementedFeature until switched on;
  * with nothing to do (divergence 0, no lock) the extension path equals the plain path;
  * the divergence arithmetic as stated in csrc/objects_ext_kernels.cu, with libear's OWN extent
    panner (PolarExtent::handle on the three Cartesian source positions, oracle/_ref) doing the
    panning: three sources, power weights (1 - x) / (1 + x), x / (1 + x), x / (1 + x), combined as
    sqrt(sum w g^2);
  * channel lock: an object within maxDistance of a loudspeaker renders as if it sat on it, one
    outside is untouched; power is preserved by both features."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
LAYOUT = "4+7+0"


def cart(az, el, d):
    az, el = np.radians(az), np.radians(el)
    return np.stack([np.sin(-az) * np.cos(el), np.cos(-az) * np.cos(el), np.sin(el)], -1) * np.asarray(d)[..., None]


def run_ext(eb, gc, cols, div=None, rng_=None, lock=None):
    import torch
    dev = torch.device("cuda", 0)
    n = len(cols[0])
    d_in = torch.from_numpy(np.stack(cols)).to(dev)
    M = gc.num_outputs
    d = torch.full((n, M), float("nan"), dtype=torch.float32, device=dev)
    f = torch.full((n, M), float("nan"), dtype=torch.float32, device=dev)
    keep = [None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).to(dev)
            for a in (div, rng_, lock)]
    gc.calculate_ext_device(n, [d_in[k].data_ptr() for k in range(8)], d.data_ptr(), f.data_ptr(),
                            *[None if t is None else t.data_ptr() for t in keep],
                            stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return d.cpu().numpy(), f.cpu().numpy()


def objects(n, seed, extent=True):
    rng = np.random.default_rng(seed)
    w = rng.uniform(0, 120, n) if extent else np.zeros(n)
    h = rng.uniform(0, 90, n) if extent else np.zeros(n)
    return [rng.uniform(-180, 180, n), rng.uniform(-60, 60, n), np.ones(n), w, h, np.zeros(n),
            rng.uniform(0.2, 1, n), rng.uniform(0, 1, n)]


def test_default_is_libears_not_implemented(eb):
    gc = eb.GainCalculatorObjects(LAYOUT)
    cols = objects(16, 1)
    with pytest.raises(eb.NotImplementedFeature, match="divergence"):
        run_ext(eb, gc, cols, div=np.full(16, 0.5))
    with pytest.raises(eb.NotImplementedFeature, match="channelLock"):
        run_ext(eb, gc, cols, lock=np.full(16, 0.3))
    gc.enable_extensions(divergence=True)
    with pytest.raises(eb.NotImplementedFeature, match="channelLock"):
        run_ext(eb, gc, cols, lock=np.full(16, 0.3))


def test_nothing_to_do_equals_the_plain_path(eb):
    gc = eb.GainCalculatorObjects(LAYOUT)
    gc.enable_extensions(divergence=True, channel_lock=True)
    cols = objects(3000, 2)
    d0, f0 = gc.calculate(*cols)
    d, f = run_ext(eb, gc, cols, div=np.zeros(3000), lock=np.full(3000, -1.0))
    assert np.abs(d - d0).max() < 1e-7 and np.abs(f - f0).max() < 1e-7


def test_divergence_against_the_stated_arithmetic_over_libears_extent_panner(eb, ref):
    n = 2000
    cols = objects(n, 3)
    rng = np.random.default_rng(4)
    x = rng.uniform(0, 1, n)
    x[:50] = 1.0
    r = rng.uniform(10, 90, n)
    gc = eb.GainCalculatorObjects(LAYOUT)
    gc.enable_extensions(divergence=True)
    d, f = run_ext(eb, gc, cols, div=x, rng_=r)
    az, el = cols[0], cols[1]
    # rows of the local frame: right, front, up (ref include/ear/common/geom.hpp:91-98)
    M = np.stack([cart(az - 90, np.zeros(n), np.ones(n)), cart(az, el, np.ones(n)), cart(az, el + 90, np.ones(n))], 1)
    pos = [cart(az, el, np.ones(n))]
    for sgn in (1, -1):
        loc = cart(sgn * x * r, np.zeros(n), np.ones(n))
        pos.append(np.einsum("nk,nkc->nc", loc, M))
    wts = [(1 - x) / (1 + x), x / (1 + x), x / (1 + x)]
    ext = ref.Extent(LAYOUT)
    acc = 0
    for p, w in zip(pos, wts):
        g = ext.handle(p, cols[3], cols[4], cols[5])
        acc = acc + w[:, None] * g ** 2
    g = np.sqrt(acc) * cols[6][:, None]
    lfe = eb.Layout(LAYOUT).is_lfe
    exp_d = np.zeros((n, len(lfe)))
    exp_d[:, ~lfe] = g * np.sqrt(1 - cols[7])[:, None]
    exp_f = np.zeros_like(exp_d)
    exp_f[:, ~lfe] = g * np.sqrt(cols[7])[:, None]
    assert np.abs(d - exp_d).max() < 1e-5 and np.abs(f - exp_f).max() < 1e-5
    # power is preserved: |direct|^2 + |diffuse|^2 = gain^2
    tot = np.sqrt((d.astype(np.float64) ** 2).sum(1) + (f.astype(np.float64) ** 2).sum(1))
    assert np.abs(tot - cols[6]).max() < 1e-5
    # full divergence: nothing left at the centre position for a point source straight ahead
    one = [np.zeros(1), np.zeros(1), np.ones(1), np.zeros(1), np.zeros(1), np.zeros(1), np.ones(1), np.zeros(1)]
    d1, _ = run_ext(eb, gc, one, div=np.ones(1), rng_=np.full(1, 30.0))
    names = eb.Layout(LAYOUT).channel_names
    assert abs(d1[0, names.index("M+030")] - np.sqrt(0.5)) < 1e-6 and abs(d1[0, names.index("M-030")] - np.sqrt(0.5)) < 1e-6
    assert abs(d1[0, names.index("M+000")]) < 1e-6


def test_channel_lock_snaps_to_the_nearest_loudspeaker_within_range(eb):
    lay = eb.Layout(LAYOUT)
    keep = ~lay.is_lfe
    spk = lay.positions[keep]                      # az, el, distance
    rng = np.random.default_rng(6)
    n = len(spk)
    near = [spk[:, 0] + rng.uniform(-3, 3, n), spk[:, 1] + rng.uniform(-3, 3, n), np.ones(n), np.zeros(n), np.zeros(n),
            np.zeros(n), np.ones(n), np.zeros(n)]
    gc = eb.GainCalculatorObjects(LAYOUT)
    gc.enable_extensions(channel_lock=True)
    d, _ = run_ext(eb, gc, near, lock=np.full(n, 0.2))
    one_hot = np.zeros((n, len(lay.is_lfe)))
    one_hot[np.arange(n), np.where(keep)[0]] = 1.0
    assert np.abs(d - one_hot).max() < 1e-6
    # out of range: untouched; flag off (negative): untouched; no limit (+inf): always snaps
    plain, _ = gc.calculate(*near)
    d_far, _ = run_ext(eb, gc, near, lock=np.full(n, 1e-3))
    d_off, _ = run_ext(eb, gc, near, lock=np.full(n, -1.0))
    assert np.abs(d_far - plain).max() < 1e-7 and np.abs(d_off - plain).max() < 1e-7
    far = objects(500, 7, extent=False)
    d_inf, _ = run_ext(eb, gc, far, lock=np.full(500, np.inf))
    assert (np.sort(d_inf / far[6][:, None] / np.sqrt(1 - far[7])[:, None], axis=1)[:, -1] > 1 - 1e-6).all()

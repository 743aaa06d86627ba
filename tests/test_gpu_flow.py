"""The Objects signal flow (ref docs/dsp.rst:40-71) as one device pipeline, against
the same chain composed from the compiled reference's own classes: two GainInterpolators, one
BlockConvolver per loudspeaker channel with designDecorrelators filters, DelayBuffer(255), sum.
Gains come from the Objects calculator (moving extent objects).  Tolerance: 1e-6 of peak."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def reference_flow(ref, layout, x, samples, direct, diffuse):
    n_out = direct.shape[2]
    bus_direct = ref.interp_matrix(x, samples, direct)
    bus_diffuse = ref.interp_matrix(x, samples, diffuse)
    filt = ref.decorrelators(layout)
    out = ref.delay_process(bus_direct, ref.decorrelator_delay())
    for c in range(n_out):
        conv = ref.Convolver(512, 1)
        fid, _ = conv.add_filter(filt[c])
        conv.set_filter(fid)
        out[c] += conv.process(bus_diffuse[c])
    return out


@pytest.mark.parametrize("layout,n_obj", [("0+5+0", 8), ("9+10+3", 5), ("4+7+0", 40)])
def test_objects_flow_matches_reference_chain(eb, ref, layout, n_obj):
    rng = np.random.default_rng(31)
    T = 512 * 12
    x = np.clip(rng.normal(0, 0.25, (n_obj, T)), -1, 1).astype(np.float32)
    samples = np.array([0, 1000, 1000, 3072, 5000, 7000])
    gc = eb.GainCalculatorObjects(layout)
    az0, el = rng.uniform(-180, 180, n_obj), rng.uniform(-30, 60, n_obj)
    w, h = rng.uniform(0, 90, n_obj), rng.uniform(0, 40, n_obj)
    dif = rng.uniform(0, 1, n_obj)
    direct, diffuse = [], []
    for k in range(len(samples)):
        d, f = gc.calculate((az0 + 20.0 * k + 180) % 360 - 180, el, 1.0, w, h, 0.0, 1.0, dif)
        direct.append(d)
        diffuse.append(f)
    direct, diffuse = np.stack(direct), np.stack(diffuse)
    r = eb.ObjectsRenderer(layout, n_obj)
    assert r.block_size == 512 and r.delay == 255
    r.set_gains(samples, direct, diffuse)
    # two calls: the convolver tails, the delay memory and the interpolation position carry over
    out = np.concatenate([r.process(0, x[:, :512 * 5]), r.process(512 * 5, x[:, 512 * 5:])], axis=1)
    exp = reference_flow(ref, layout, x, samples, direct, diffuse)
    assert np.abs(out - exp).max() / np.abs(exp).max() < 1e-6


def test_objects_flow_rejects_partial_blocks(eb):
    r = eb.ObjectsRenderer("0+5+0", 2)
    with pytest.raises(eb.InvalidArgument):
        r.process(0, np.zeros((2, 700), np.float32))

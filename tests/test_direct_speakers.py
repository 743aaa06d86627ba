"""GainCalculatorDirectSpeakers against the golden vectors of the compiled
libear reference (tests/golden/direct_speakers.npz: the cases of the reference's
tests/gain_calculator_direct_speakers_tests.cpp plus a seeded sweep over all layouts).

Routing (labels, mapping rules, bounds, LFE) is host-side table work and is checked on CPU;
channels that fall through to the point-source panner need the device (K1) and are checked
under -m gpu.  Gains are FP64 on both sides: held to 1e-12, one-hot routes bit-exact."""
import json
import os

import numpy as np
import pytest

GOLDEN = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "direct_speakers.npz"))
KW = ("az", "el", "dist", "bounds", "low_pass", "high_pass", "pack", "cartesian", "screen_edge_lock")


def cases():
    for i in range(int(GOLDEN["n"])):
        yield (json.loads(str(GOLDEN[f"{i}/case"])), GOLDEN[f"{i}/gains"], list(GOLDEN[f"{i}/warnings"]),
               int(GOLDEN[f"{i}/error"]))


def run_all(eb, need_device):
    calcs, n_checked, n_panned = {}, 0, 0
    for c, gains, warnings, error in cases():
        key = (c["layout"], json.dumps(c.get("substitutions")))
        if key not in calcs:
            calcs[key] = eb.GainCalculatorDirectSpeakers(c["layout"], c.get("substitutions"))
        kw = {k: c[k] for k in KW if k in c}
        try:
            out, w = calcs[key].calculate(c.get("labels", []), **kw)
        except eb.CudaError:
            assert not need_device, "no CUDA device on a GPU run"
            assert error == 0, c   # only the point-source fallback may ask for the device
            n_panned += 1
            continue
        except eb.NotImplementedFeature:
            assert error == 1, c
            n_checked += 1
            continue
        except eb.AdmError:
            assert error == 3, c
            n_checked += 1
            continue
        assert error == 0, c
        assert w == warnings, c
        assert np.abs(out - gains).max() < 1e-12, c
        n_checked += 1
    return n_checked, n_panned


def test_routing_matches_reference_on_host(eb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("covered by the gpu test")
    n_checked, n_panned = run_all(eb, need_device=False)
    assert n_checked > 300 and n_panned > 50


def test_live_against_compiled_reference(eb, ref):
    """Fresh random labels / packs straight against oracle/_ref (host-routed cases)."""
    rng = np.random.default_rng(5)
    for layout in ("9+10+3", "4+5+1", "0+7+0"):
        names = ref.layout_info(layout)["names"]
        gc = eb.GainCalculatorDirectSpeakers(layout)
        for _ in range(60):
            label = names[rng.integers(len(names))]
            pack = ["AP_00010009", "AP_00010010", None][rng.integers(3)]
            exp, codes = ref.direct_speakers_calc(layout, [label], pack_id=pack, codes=True)
            out, w = gc.calculate([label], pack=pack)
            assert w == codes and np.abs(out - exp).max() == 0.0


@pytest.mark.gpu
def test_all_cases_on_device(eb):
    n_checked, n_panned = run_all(eb, need_device=True)
    assert n_checked == int(GOLDEN["n"]) and n_panned == 0


@pytest.mark.gpu
def test_batch_is_one_launch(eb):
    """A whole bed in one call: the panned channels share one K1 launch."""
    gc = eb.GainCalculatorDirectSpeakers("4+7+0")
    chans = [dict(labels=["foo"], az=float(a), el=5.0) for a in np.linspace(-170, 170, 64)]
    chans += [dict(labels=["M+030"], az=30.0), dict(labels=["LFE1"], az=45.0, el=-30.0)]
    l0 = gc.launch_count
    g, w = gc.calculate_batch(chans)
    assert gc.launch_count - l0 == 1
    single = np.stack([gc.calculate(**c)[0] for c in chans])
    assert np.abs(g - single).max() == 0.0
    assert np.allclose((g[:64] ** 2).sum(1), 1.0, atol=1e-12)


def test_invalid_layouts_fail_at_construction_like_libear():
    """libear builds the point-source panner in the GainCalculatorDirectSpeakers constructor
    (ref src/direct_speakers/gain_calculator_direct_speakers.cpp:84), so a layout it cannot pan on
    fails THERE, not at the first channel that happens to need panning.  No device needed: the
    region tables are validated on the host."""
    import libear_b200 as eb
    names = ["M+000", "M+030", "M-030", "M+110", "M-110"]
    pos = [[0, 0, 1], [0, 0, 1], [-30, 0, 1], [110, 0, 1], [-110, 0, 1]]   # two loudspeakers at one place
    with pytest.raises(eb.InternalError):
        eb.GainCalculatorDirectSpeakers(eb.Layout("broken", channels=dict(names=names, pos=pos)))

"""CUDA path (through the C ABI) against the committed golden vectors of the compiled libear
reference (tests/golden/*.npz) -- needs no oracle library on the GPU box.

Tolerances (north star): region indices bit-exact, gains within 1e-5 absolute (FP64 paths are
held to 1e-12), rendered audio within 1e-6 of peak.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LAYOUTS = ["0+2+0", "0+5+0", "2+5+0", "4+5+0", "4+5+1", "3+7+0", "4+9+0", "9+10+3", "0+7+0", "4+7+0"]


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.mark.parametrize("layout", LAYOUTS)
def test_psp(eb, layout):
    g = golden("psp")
    gains, reg, inner = eb.PointSourcePanner(layout).handle(g[f"{layout}/xyz"], regions=True)
    assert np.abs(gains - g[f"{layout}/gains"]).max() < 1e-12
    if layout == "0+2+0":
        # the stereo downmix has no region list of its own (the reference probe reports -1);
        # ours reports the region of the 0+5+0 panner it pans on (point_source_panner.cpp:369-392)
        g5 = golden("psp")
        _, reg5, inner5 = eb.PointSourcePanner("0+5+0").handle(g[f"{layout}/xyz"], regions=True)
        assert (reg == reg5).all() and (inner == inner5).all()
        return
    assert (reg == g[f"{layout}/region"]).all()
    assert (inner == g[f"{layout}/inner"]).all()


@pytest.mark.parametrize("layout", ["9+10+3", "0+5+0", "4+7+0"])
def test_extent(eb, layout):
    g = golden("extent")
    gc = eb.GainCalculatorObjects(layout)
    xyz, w, h, d = (g[f"{layout}/{k}"] for k in ("xyz", "width", "height", "depth"))
    out = gc.extent(xyz, w, h, d)
    assert np.abs(out - g[f"{layout}/simd"]).max() < 1e-5
    assert np.abs(out - g[f"{layout}/scalar"]).max() < 1e-5
    unit = xyz / np.linalg.norm(xyz, axis=1, keepdims=True)
    assert np.abs(gc.extent(unit, w, h, 0.0) - g[f"{layout}/double_nodepth_unit"]).max() < 1e-5


@pytest.mark.parametrize("layout", ["0+5+0", "9+10+3", "4+7+0", "0+2+0"])
def test_objects(eb, layout):
    g = golden("objects")
    gc = eb.GainCalculatorObjects(layout)
    args = list(g[f"{layout}/in"])
    d32, f32, reg = gc.calculate(*args, dtype=np.float32, regions=True)
    assert np.abs(d32 - g[f"{layout}/direct32"]).max() < 1e-5
    assert np.abs(f32 - g[f"{layout}/diffuse32"]).max() < 1e-5
    if layout != "0+2+0":
        assert (reg == g[f"{layout}/region"]).all()
    d64, f64 = gc.calculate(*args, dtype=np.float64)
    tol = 1e-12 if layout == "0+5+0" else 1e-5
    assert np.abs(d64 - g[f"{layout}/direct64"]).max() < tol
    assert np.abs(f64 - g[f"{layout}/diffuse64"]).max() < tol


def test_interp(eb):
    g = golden("interp")
    x, s, gains = g["matrix/x"], g["matrix/samples"], g["matrix/gains"]
    peak = np.abs(g["matrix/out"]).max()
    gi = eb.GainInterpolator(x.shape[0], gains.shape[2])
    gi.set_points(s, gains)
    assert np.abs(gi.process(0, x) - g["matrix/out"]).max() / peak < 1e-6
    assert np.abs(gi.process(250, x[:, :1000]) - g["matrix/out_start_250"]).max() / peak < 1e-6
    # the same buffer in calls of 75 samples (find_block caching, segment boundaries)
    parts = [gi.process(t, x[:, t:t + 75]) for t in range(0, x.shape[1], 75)]
    assert np.abs(np.concatenate(parts, axis=1) - g["matrix/out"]).max() / peak < 1e-6
    gv = eb.GainInterpolator(1, gains.shape[2])
    gv.set_points(s, gains[:, :1, :])
    assert np.abs(gv.process(0, x[:1]) - g["vector/out"]).max() / np.abs(g["vector/out"]).max() < 1e-6
    gs = eb.GainInterpolator(1, 1)
    gs.set_points(s, gains[:, :1, :1])
    assert np.abs(gs.process(0, x[:1])[0] - g["single/out"]).max() / np.abs(g["single/out"]).max() < 1e-6


def test_interp_cfg3(eb):
    g = golden("interp")
    x = g["cfg3/x16"].astype(np.float32)
    gi = eb.GainInterpolator(x.shape[0], g["cfg3/gains"].shape[2])
    gi.set_points(g["cfg3/samples"], g["cfg3/gains"])
    out = gi.process(0, x)
    assert np.abs(out - g["cfg3/out"]).max() / np.abs(g["cfg3/out"]).max() < 1e-6


def test_convolver(eb):
    g = golden("convolver")
    B, nb = int(g["B"]), int(g["nb"])
    x = g["x"][None, :]
    c = eb.BlockConvolver(B, 1, nb)
    a, na = c.add_filter(g["f0"])
    b, nbl = c.add_filter(g["f1"])
    assert (na, nbl) == (4, 2)
    c.set_filter(a)
    y = [c.process(x[:, :3 * B])]
    c.crossfade_filter(b)
    y.append(c.process(x[:, 3 * B:7 * B]))
    c.crossfade_filter(None)
    y.append(c.process(x[:, 7 * B:9 * B]))
    y.append(c.process(None, 2))
    c.crossfade_filter(a)
    y.append(c.process(x[:, 11 * B:]))
    y = np.concatenate(y, axis=1)[0]
    assert np.abs(y - g["y"]).max() < 1e-6 * max(1.0, np.abs(g["y"]).max())


def test_convolver_cfg4(eb):
    g = golden("convolver")
    dec = eb.design_decorrelators("9+10+3")
    x, chans = g["cfg4/x"], g["cfg4/channels"]
    c = eb.BlockConvolver(512, len(chans), 1)
    for i, ch in enumerate(chans):
        fid, _ = c.add_filter(dec[ch])
        c.set_filter(fid, channel=i)
    y = c.process(x)
    assert np.abs(y - g["cfg4/y"]).max() / np.abs(g["cfg4/y"]).max() < 1e-6


@pytest.mark.parametrize("layout,norm", [("4+7+0", "SN3D"), ("4+7+0", "N3D"), ("4+7+0", "FuMa"),
                                         ("9+10+3", "SN3D"), ("0+5+0", "N3D")])
def test_hoa(eb, layout, norm):
    g = golden("hoa")
    D, warnings = eb.GainCalculatorHOA(layout).calculate(g["orders"], g["degrees"], norm)
    assert warnings == []
    assert np.abs(D - g[f"{layout}/{norm}"]).max() < 1e-9

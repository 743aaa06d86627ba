"""Pins the oracle: the numpy restatement (oracle/ear_oracle.py) against the golden vectors
produced by the compiled, unmodified libear reference (tests/golden/*.npz, generator
tests/golden/generate.py), against the known answers the reference's own tests hold
(tests/golden/kat.npz), and -- where oracle/_ref is built -- against the reference live.

CPU only.  Tolerances: FP64 paths 1e-12; paths through the float32 extent core 2e-6 (the
reference's own SIMD and scalar cores differ by 4.6e-7); audio 1e-6 of
peak.  Region indices must be identical.
"""
import json
import os

import numpy as np
import pytest

from oracle import ear_oracle as eo

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LAYOUTS = ["0+2+0", "0+5+0", "2+5+0", "4+5+0", "4+5+1", "3+7+0", "4+9+0", "9+10+3", "0+7+0", "4+7+0"]


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


# ------------------------------------------------------------------ point source panner
@pytest.mark.parametrize("layout", LAYOUTS)
def test_psp_golden(layout):
    g = golden("psp")
    gains, reg, inner = eo.PointSourcePanner(layout).handle(g[f"{layout}/xyz"])
    assert np.abs(gains - g[f"{layout}/gains"]).max() < 1e-12
    if layout != "0+2+0":  # the stereo downmix has no region list of its own
        assert (reg == g[f"{layout}/region"]).all()
        assert (inner == g[f"{layout}/inner"]).all()


def test_psp_known_answers():
    # tests/point_source_panner_tests.cpp:421-519: one-hot at every loudspeaker; +-15 deg on 0+5+0
    kat = golden("kat")
    p = eo.PointSourcePanner("0+5+0")
    assert np.abs(p.handle(eo.cart(15.0, 0.0))[0][0] - kat["psp_0+5+0_az15"]).max() < 1e-12
    for layout in LAYOUTS[1:]:
        lay = eo.Layout(layout)
        pos = lay.pos[~lay.is_lfe]
        g = eo.PointSourcePanner(layout).handle(eo.cart(pos[:, 0], pos[:, 1]))[0]
        assert np.abs(g - np.eye(len(pos))).max() < 1e-9


@pytest.mark.parametrize("layout", ["0+5+0", "9+10+3", "4+7+0"])
def test_psp_properties(layout):
    # tests/point_source_panner_tests.cpp:341-419: non-negative, unit norm
    rng = np.random.default_rng(5)
    v = rng.normal(size=(200, 3))
    g = eo.PointSourcePanner(layout).handle(v)[0]
    assert g.min() >= 0.0
    assert np.abs(np.linalg.norm(g, axis=1) - 1.0).max() < 1e-12


# ------------------------------------------------------------------ extent / objects
@pytest.mark.parametrize("layout", ["9+10+3", "0+5+0", "4+7+0"])
def test_extent_golden(layout):
    g = golden("extent")
    ext = eo.PolarExtent(layout)
    xyz, w, h, d = (g[f"{layout}/{k}"] for k in ("xyz", "width", "height", "depth"))
    n = 40 if layout == "9+10+3" else 16
    out = np.array([ext.handle(xyz[i], w[i], h[i], d[i]) for i in range(n)])
    assert np.abs(out - g[f"{layout}/scalar"][:n]).max() < 2e-6
    assert np.abs(out - g[f"{layout}/simd"][:n]).max() < 2e-6
    # the reference's float cores against its own double-precision restatement
    # (tests/extent_tests.cpp:136-166 uses isApprox(1e-5))
    assert np.abs(g[f"{layout}/simd_nodepth_unit"] - g[f"{layout}/double_nodepth_unit"]).max() < 1e-5


def test_extent_zero_is_point_source():
    # tests/extent_tests.cpp:110-134: zero extent at distance 1 equals the point-source panner
    ext = eo.PolarExtent("9+10+3")
    rng = np.random.default_rng(6)
    for v in rng.normal(size=(10, 3)):
        v /= np.linalg.norm(v)
        assert np.abs(ext.handle(v, 0.0, 0.0, 0.0) - ext.psp.handle_one(v)[0]).max() < 1e-12


@pytest.mark.parametrize("layout", ["0+5+0", "9+10+3", "4+7+0"])
def test_objects_golden(layout):
    g = golden("objects")
    calc = eo.GainCalculatorObjects(layout)
    args = g[f"{layout}/in"]
    n = 32 if layout != "0+5+0" else 128
    tol = 1e-12 if layout == "0+5+0" else 2e-6
    for i in range(n):
        d, f = calc.calculate(*args[:, i])
        assert np.abs(d - g[f"{layout}/direct64"][i]).max() < tol
        assert np.abs(f - g[f"{layout}/diffuse64"][i]).max() < tol
        assert np.abs(d.astype(np.float32) - g[f"{layout}/direct32"][i]).max() < max(tol, 1e-7)


# ------------------------------------------------------------------ gain interpolator
def test_interp_golden():
    g = golden("interp")
    x, s, gains = g["matrix/x"], g["matrix/samples"], g["matrix/gains"]
    out = eo.interp_matrix(x, s, gains)
    peak = np.abs(g["matrix/out"]).max()
    assert np.abs(out - g["matrix/out"]).max() / peak < 1e-6
    # splitting the buffer into calls of 75 samples does not change the reference's result
    assert np.array_equal(g["matrix/out"], g["matrix/out_calls_of_75"])
    out250 = eo.interp_matrix(x[:, :1000], s, gains, block_start=250)
    assert np.abs(out250 - g["matrix/out_start_250"]).max() / peak < 1e-6
    # vector / single policies are the N = 1 (and M = 1) cases of the matrix policy
    v = eo.interp_matrix(x[:1], s, gains[:, :1, :])
    assert np.abs(v - g["vector/out"]).max() / np.abs(g["vector/out"]).max() < 1e-6
    sg = eo.interp_matrix(x[:1], s, gains[:, :1, :1])
    assert np.abs(sg[0] - g["single/out"]).max() / np.abs(g["single/out"]).max() < 1e-6


def test_interp_golden_cfg3():
    g = golden("interp")
    out = eo.interp_matrix(g["cfg3/x16"].astype(np.float32), g["cfg3/samples"], g["cfg3/gains"])
    assert np.abs(out - g["cfg3/out"]).max() / np.abs(g["cfg3/out"]).max() < 1e-6


def test_interp_closed_form():
    # tests/gain_interpolator_tests.cpp:58-81: single ramp against the closed form
    x = np.ones((1, 100), dtype=np.float32)
    out = eo.interp_matrix(x, [20, 60], np.array([[[0.0]], [[1.0]]], dtype=np.float32))[0]
    exp = np.clip((np.arange(100) - 20) / 40.0, 0, 1)
    assert np.abs(out - exp).max() < 1e-6


# ------------------------------------------------------------------ block convolver
def test_convolver_golden():
    g = golden("convolver")
    B, nb = int(g["B"]), int(g["nb"])
    x = g["x"]
    c = eo.BlockConvolver(B, nb)
    a, b = c.add_filter(g["f0"]), c.add_filter(g["f1"])
    c.set_filter(a)
    y = [c.process(x[:3 * B])]
    c.crossfade_filter(b)
    y.append(c.process(x[3 * B:7 * B]))
    c.crossfade_filter(None)
    y.append(c.process(x[7 * B:9 * B]))
    y.append(c.process(None, 2))
    c.crossfade_filter(a)
    y.append(c.process(x[11 * B:]))
    y = np.concatenate(y)
    assert np.abs(y - g["y"]).max() < 1e-6 * max(1.0, np.abs(g["y"]).max())


def test_convolver_golden_cfg4():
    g = golden("convolver")
    dec = golden("decorrelators")["9+10+3/f32"]
    for x, ch, y_ref in zip(g["cfg4/x"], g["cfg4/channels"], g["cfg4/y"]):
        c = eo.BlockConvolver(512, 1)
        c.set_filter(c.add_filter(dec[ch]))
        y = c.process(x)
        assert np.abs(y - y_ref).max() / np.abs(y_ref).max() < 1e-6


def test_convolver_vs_direct_convolution():
    # tests/block_convolver_tests.cpp:29-195: against an O(N L) direct convolution
    rng = np.random.default_rng(7)
    x, f = rng.normal(size=64 * 6), rng.normal(size=150)
    c = eo.BlockConvolver(64, 3)
    c.set_filter(c.add_filter(f))
    assert np.abs(c.process(x) - np.convolve(x, f)[:len(x)]).max() < 1e-9


# ------------------------------------------------------------------ decorrelators
def test_decorrelator_known_answers():
    kat = golden("kat")  # tests/decorrelate_tests.cpp:19-28
    d = eo.design_decorrelator_basic(7)
    assert len(d) == 512
    assert np.abs(d[kat["decorrelator7_taps"]] - kat["decorrelator7_values"]).max() < 1e-12


@pytest.mark.parametrize("layout", ["0+5+0", "4+7+0", "9+10+3"])
def test_decorrelators_golden(layout):
    g = golden("decorrelators")
    d = eo.design_decorrelators(layout)
    assert np.abs(d - g[f"{layout}/f64"]).max() < 1e-12
    assert np.abs(d.astype(np.float32) - g[f"{layout}/f32"]).max() < 1e-7
    assert int(g["delay"]) == 255


# ------------------------------------------------------------------ HOA
@pytest.mark.parametrize("layout,norm", [("4+7+0", "SN3D"), ("4+7+0", "N3D"), ("4+7+0", "FuMa"),
                                         ("9+10+3", "SN3D"), ("0+5+0", "N3D")])
def test_hoa_golden(layout, norm):
    g = golden("hoa")
    D = eo.hoa_decode_matrix(layout, g["orders"], g["degrees"], norm)
    assert np.abs(D - g[f"{layout}/{norm}"]).max() < 1e-10


def test_hoa_points_known_answers():
    kat, g = golden("kat"), golden("hoa")  # tests/hoa_tests.cpp:7-16
    assert int(g["n_points"]) == int(kat["hoa_n_points"]) == 5200
    assert np.abs(g["points_first_last"][0] - kat["hoa_point_first"]).max() < 1e-9
    assert np.abs(g["points_first_last"][1] - kat["hoa_point_last"]).max() < 1e-9
    pts = np.load(os.path.join(os.path.dirname(eo.__file__), "data", "tdesign_5200.npy"))
    assert np.abs(pts[[0, -1]] - g["points_first_last"]).max() < 1e-12


# ------------------------------------------------------------------ live reference
def test_restatement_vs_live_reference(ref):
    """Where oracle/_ref is built: fresh random inputs through both."""
    rng = np.random.default_rng(int.from_bytes(os.urandom(4), "little"))
    for layout in ("0+5+0", "4+9+0", "9+10+3"):
        v = rng.normal(size=(100, 3))
        a, ra, ia = eo.PointSourcePanner(layout).handle(v)
        b, rb, ib = ref.PointSourcePanner(layout).handle(v)
        assert np.abs(a - b).max() < 1e-12 and (ra == rb).all() and (ia == ib).all()
    ext = eo.PolarExtent("9+10+3")
    rext = ref.Extent("9+10+3", 1)
    v = rng.normal(size=(20, 3))
    w, h = rng.uniform(0, 360, 20), rng.uniform(0, 360, 20)
    out = np.array([ext.handle(v[i], w[i], h[i], 0.3) for i in range(20)])
    assert np.abs(out - rext.handle(v, w, h, 0.3)).max() < 2e-6


def test_golden_fixtures_are_current(ref):
    """The committed fixtures are what the compiled reference produces today."""
    g = golden("psp")
    for layout in ("0+5+0", "9+10+3"):
        gains, reg, _ = ref.PointSourcePanner(layout).handle(g[f"{layout}/xyz"])
        assert np.array_equal(gains, g[f"{layout}/gains"]) and np.array_equal(reg, g[f"{layout}/region"])
    g = golden("direct_speakers")
    for i in range(int(g["n"])):
        c = json.loads(str(g[f"{i}/case"]))
        if int(g[f"{i}/error"]):
            with pytest.raises(ref.RefError):
                ref.direct_speakers_calc(c["layout"], c.get("labels", []), pack_id=c.get("pack"),
                                         cartesian=c.get("cartesian", False),
                                         screen_edge_lock=c.get("screen_edge_lock", 0))
            continue
        gains, codes = ref.direct_speakers_calc(
            c["layout"], c.get("labels", []), c.get("az", 0.0), c.get("el", 0.0), c.get("dist", 1.0),
            c.get("bounds"), c.get("low_pass"), c.get("high_pass"), c.get("pack"), c.get("substitutions"),
            screen_edge_lock=c.get("screen_edge_lock", 0), codes=True)
        assert np.abs(gains - g[f"{i}/gains"]).max() < 1e-15 and codes == list(g[f"{i}/warnings"])


# ------------------------------------------------------------------ Cartesian <-> polar conversion
def _oracle_conversion(pos, ext, to_polar):
    fn = eo.extent_cart_to_polar if to_polar else eo.extent_polar_to_cart
    out = [fn(p, e) for p, e in zip(pos, ext)]
    return np.array([o[0] for o in out]), np.array([o[1] for o in out])


def test_conversion_golden():
    g = golden("conversion")
    pos, ext = _oracle_conversion(g["polar"], g["polar_ext"], False)
    assert np.abs(pos - g["p2c_pos"]).max() < 1e-12
    assert np.abs(ext - g["p2c_ext"]).max() < 1e-12
    pos, ext = _oracle_conversion(g["cart"], g["cart_ext"], True)
    assert np.abs(pos - g["c2p_pos"]).max() < 1e-10  # degrees
    assert np.abs(ext - g["c2p_ext"]).max() < 1e-10


def test_conversion_known_answers():
    # tests/conversion_tests.cpp:162-185 (margin 1e-6 there)
    kat = golden("kat")
    pos, ext = eo.extent_polar_to_cart([10.0, 20.0, 0.3], [40.0, 50.0, 0.6])
    assert np.abs(np.concatenate([pos, ext]) - kat["conversion_p2c"]).max() < 1e-9
    pos, ext = eo.extent_cart_to_polar([0.9, 0.8, 0.1], [0.3, 0.5, 0.4])
    assert np.abs(np.concatenate([pos, ext]) - kat["conversion_c2p"]).max() < 1e-9


def test_conversion_round_trip_and_poles():
    # tests/conversion_tests.cpp:52-90
    for az in (0.0, -10.0, 10.0, 90.0, -90.0, 150.0, -150.0):
        for el in (0.0, -10.0, 10.0, -45.0, 45.0, -90.0, 90.0):
            for dist in (0.5, 1.0):
                back = eo.point_cart_to_polar(*eo.point_polar_to_cart(az, el, dist))
                if abs(el) < 90.0:
                    assert abs(back[0] - az) < 1e-6
                assert abs(back[1] - el) < 1e-6 and abs(back[2] - dist) < 1e-6
    for sign in (-1.0, 1.0):
        for d in (0.5, 1.0, 2.0):
            assert np.abs(eo.point_polar_to_cart(0.0, sign * 90.0, d) - [0.0, 0.0, sign * d]).max() < 1e-6
            assert np.abs(eo.point_cart_to_polar(0.0, 0.0, sign * d) - [0.0, sign * 90.0, d]).max() < 1e-6


def test_conversion_live(ref):
    rng = np.random.default_rng(23)
    n = 300
    cartp, ext = rng.uniform(-1.5, 1.5, (n, 3)), rng.uniform(0, 1, (n, 3))
    r_pos, r_ext = ref.conversion(cartp, ext, to_polar=True)
    pos, e = _oracle_conversion(cartp, ext, True)
    assert np.abs(pos - r_pos).max() < 1e-10 and np.abs(e - r_ext).max() < 1e-10


# ------------------------------------------------------------------ custom layouts / convex hull
def _golden_facets(g, k):
    flat, out, i = g[f"hull{k}/facets"], [], 0
    while i < len(flat):
        out.append([int(v) for v in flat[i + 1:i + 1 + flat[i]]])
        i += 1 + flat[i]
    return out


def test_convex_hull_golden():
    g = golden("custom")
    for k in range(int(g["n_clouds"])):
        assert eo.convex_hull(g[f"hull{k}/xyz"]) == _golden_facets(g, k)


def _custom_layout(g, key):
    return dict(name=key, names=[str(n) for n in g[f"{key}/names"]], pos=g[f"{key}/pos"],
                nominal=g[f"{key}/nominal"], az_range=g[f"{key}/az_range"], el_range=g[f"{key}/el_range"],
                is_lfe=g[f"{key}/is_lfe"])


def test_custom_layout_psp_golden():
    g = golden("custom")
    for key in g["layouts"]:
        key = str(key)
        gains, reg, inner = eo.PointSourcePanner(_custom_layout(g, key)).handle(g[f"{key}/xyz"])
        assert np.array_equal(reg, g[f"{key}/region"]), key
        assert np.array_equal(inner, g[f"{key}/inner"]), key
        assert np.abs(gains - g[f"{key}/gains"]).max() < 1e-12, key

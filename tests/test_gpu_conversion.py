"""ear::conversion as a batched device pre-pass (earb200_conversion_*) against the
compiled reference's golden vectors (tests/golden/conversion.npz), the known answers of the
reference's own tests (tests/conversion_tests.cpp), and the reference live on seeded inputs.
FP64 with CUDA's libm: agreement to 1e-9 (degrees / normalised units), the reference's own
tests use 1e-6."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-9


def _close_polar(a, b, tol=TOL):
    # azimuth is meaningless at the poles and wraps at +-180
    daz = np.abs((a[:, 0] - b[:, 0] + 180.0) % 360.0 - 180.0)
    daz = np.where(np.abs(b[:, 1]) >= 90.0 - 1e-9, 0.0, daz)
    return max(daz.max(), np.abs(a[:, 1:] - b[:, 1:]).max()) < tol


def test_conversion_golden(eb):
    g = np.load(os.path.join(GOLDEN, "conversion.npz"))
    cv = eb.Conversion()
    pos, ext = cv.polar_to_cart(g["polar"], g["polar_ext"])
    assert np.abs(pos - g["p2c_pos"]).max() < TOL
    assert np.abs(ext - g["p2c_ext"]).max() < TOL
    pos, ext = cv.cart_to_polar(g["cart"], g["cart_ext"])
    assert _close_polar(pos, g["c2p_pos"])
    assert np.abs(ext - g["c2p_ext"]).max() < 1e-8
    # positions only: same positions, no extent pass
    assert np.array_equal(cv.polar_to_cart(g["polar"]), cv.polar_to_cart(g["polar"], g["polar_ext"])[0])
    assert cv.launch_count == 4


def test_conversion_known_answers(eb):
    # tests/conversion_tests.cpp:162-185
    kat = np.load(os.path.join(GOLDEN, "kat.npz"))
    cv = eb.Conversion()
    pos, ext = cv.polar_to_cart([[10.0, 20.0, 0.3]], [[40.0, 50.0, 0.6]])
    assert np.abs(np.concatenate([pos[0], ext[0]]) - kat["conversion_p2c"]).max() < 1e-9
    pos, ext = cv.cart_to_polar([[0.9, 0.8, 0.1]], [[0.3, 0.5, 0.4]])
    assert np.abs(np.concatenate([pos[0], ext[0]]) - kat["conversion_c2p"]).max() < 1e-9


def test_conversion_loop_poles_centre(eb):
    # tests/conversion_tests.cpp:52-90
    cv = eb.Conversion()
    polar = np.array([[az, el, d] for az in (0.0, -10.0, 10.0, 90.0, -90.0, 150.0, -150.0)
                      for el in (0.0, -10.0, 10.0, -45.0, 45.0, -90.0, 90.0) for d in (0.5, 1.0)])
    back = cv.cart_to_polar(cv.polar_to_cart(polar))
    assert _close_polar(back, polar, 1e-6)
    for sign in (-1.0, 1.0):
        for d in (0.5, 1.0, 2.0):
            assert np.abs(cv.polar_to_cart([[0.0, sign * 90.0, d]]) - [0.0, 0.0, sign * d]).max() < 1e-6
            assert np.abs(cv.cart_to_polar([[0.0, 0.0, sign * d]]) - [0.0, sign * 90.0, d]).max() < 1e-6
    centre = cv.polar_to_cart([[az, el, 0.0] for az in (-90.0, 0.0, 90.0) for el in (-90.0, 0.0, 90.0)])
    assert np.abs(centre).max() < 1e-6
    assert abs(cv.cart_to_polar([[0.0, 0.0, 0.0]])[0, 2]) < 1e-6


def test_conversion_whd_mapping(eb):
    # tests/conversion_tests.cpp:92-160: which Cartesian axis a polar extent axis lands on
    cv = eb.Conversion()
    for az, el, cart_whd in [(0.0, 0.0, "whd"), (90.0, 0.0, "dhw"), (-90.0, 0.0, "dhw"), (180.0, 0.0, "whd"),
                             (0.0, 90.0, "wdh"), (0.0, -90.0, "wdh")]:
        for axis_idx in range(3):
            polar_axis, cart_axis = "whd"[axis_idx], cart_whd[axis_idx]
            ext = [20.0 if polar_axis == "w" else 0.0, 20.0 if polar_axis == "h" else 0.0,
                   0.2 if polar_axis == "d" else 0.0]
            _, cext = cv.polar_to_cart([[az, el, 1.0]], [ext])
            assert "whd"[int(np.argmax(cext[0]))] == cart_axis
            cpos = cv.polar_to_cart([[az, el, 1.0]])
            _, pext = cv.cart_to_polar(cpos, [[0.1 if cart_axis == c else 0.0 for c in "whd"]])
            assert "whd"[int(np.argmax(pext[0]))] == polar_axis


def test_conversion_live_large(eb, ref):
    rng = np.random.default_rng(29)
    n = 200_000
    cv = eb.Conversion()
    cartp, ext = rng.uniform(-1.5, 1.5, (n, 3)), rng.uniform(0, 1, (n, 3))
    r_pos, r_ext = ref.conversion(cartp, ext, to_polar=True)
    pos, e = cv.cart_to_polar(cartp, ext)
    assert _close_polar(pos, r_pos)
    assert np.abs(e - r_ext).max() < 1e-8
    polar = np.stack([rng.uniform(-180, 180, n), rng.uniform(-90, 90, n), rng.uniform(0, 2, n)], 1)
    pext = np.stack([rng.uniform(0, 360, n), rng.uniform(0, 360, n), rng.uniform(0, 1, n)], 1)
    r_pos, r_ext = ref.conversion(polar, pext, to_polar=False)
    pos, e = cv.polar_to_cart(polar, pext)
    assert np.abs(pos - r_pos).max() < TOL and np.abs(e - r_ext).max() < TOL
    # round trip at full size (property, no oracle): polar -> Cartesian -> polar
    back = cv.cart_to_polar(pos)
    assert _close_polar(back, polar, 1e-6)


def test_conversion_feeds_objects_calculator(eb, ref):
    """The pre-pass in use: Cartesian-authored objects -> toPolar -> GainCalculatorObjects, against
    libear's conversion::toPolar followed by libear's calculator."""
    rng = np.random.default_rng(31)
    n = 2000
    cartp, ext = rng.uniform(-1.0, 1.0, (n, 3)), rng.uniform(0, 0.6, (n, 3))
    pos, e = eb.Conversion().cart_to_polar(cartp, ext)
    r_pos, r_e = ref.conversion(cartp, ext, to_polar=True)
    gain, diffuse = rng.uniform(0, 1, n), rng.uniform(0, 1, n)
    d, f = eb.GainCalculatorObjects("4+7+0").calculate(pos[:, 0], pos[:, 1], pos[:, 2], e[:, 0], e[:, 1], e[:, 2],
                                                      gain, diffuse)
    d_ref, f_ref = ref.Objects("4+7+0").calculate(r_pos[:, 0], r_pos[:, 1], r_pos[:, 2], r_e[:, 0], r_e[:, 1],
                                                  r_e[:, 2], gain, diffuse, dtype=np.float32)
    assert max(np.abs(d - d_ref).max(), np.abs(f - f_ref).max()) < 1e-5


def test_conversion_errors(eb):
    cv = eb.Conversion()
    assert cv.cart_to_polar(np.zeros((0, 3))).shape == (0, 3)
    with pytest.raises(eb.InternalError):  # libear: "could not find sector"
        cv.polar_to_cart([[np.nan, 0.0, 1.0]])
    # the failure does not stick to the handle
    assert np.abs(cv.polar_to_cart([[0.0, 0.0, 1.0]]) - [0.0, 1.0, 0.0]).max() < 1e-12
    with pytest.raises(eb.InternalError):
        cv.cart_to_polar([[np.nan, 1.0, 0.0]])

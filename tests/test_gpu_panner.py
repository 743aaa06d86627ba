"""K1 parity: point-source gains and region indices vs the compiled libear reference.

Mirrors tests/point_source_panner_tests.cpp of the reference (test_all_layouts :341-419,
configure_full_polar_panner :421-519) with the reference itself as the oracle.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LAYOUTS = ["0+5+0", "2+5+0", "4+5+0", "4+5+1", "3+7+0", "4+9+0", "9+10+3", "0+7+0", "4+7+0"]


def cart(az, el, d=1.0):
    az, el = np.radians(az), np.radians(el)
    return np.stack([np.sin(-az) * np.cos(el) * d, np.cos(-az) * np.cos(el) * d, np.sin(el) * d], -1)


def probe_positions(ref, layout, rng, n_random):
    info = ref.layout_info(layout)
    pos = info["pos"][~info["is_lfe"]]
    spk = cart(pos[:, 0], pos[:, 1])
    az, el = np.meshgrid(np.linspace(-180, 180, 21), np.linspace(-90, 90, 11))
    grid = cart(az.ravel(), el.ravel())
    # midpoints of all speaker pairs: lie on region edges where several regions tie
    i, j = np.triu_indices(len(spk), 1)
    mid = spk[i] + spk[j]
    mid = mid[np.linalg.norm(mid, axis=1) > 1e-6]
    rnd = rng.normal(size=(n_random, 3))
    rnd /= np.linalg.norm(rnd, axis=1, keepdims=True)
    poles = np.array([[0, 0, 1.0], [0, 0, -1.0], [0, 1.0, 0], [1e-12, 0, 1.0]])
    scaled = rnd[:1000] * rng.uniform(0.1, 3.0, size=(1000, 1))  # gains are scale invariant
    return np.concatenate([spk, grid, mid, poles, scaled, rnd])


@pytest.mark.parametrize("layout", LAYOUTS)
def test_gains_and_regions_match_reference(eb, ref, layout):
    rng = np.random.default_rng(abs(hash(layout)) % 2**32)
    xyz = probe_positions(ref, layout, rng, 100000)
    r = ref.PointSourcePanner(layout)
    g_ref, reg_ref, inner_ref = r.handle(xyz)
    p = eb.PointSourcePanner(layout)
    assert p.num_channels == r.S and p.num_regions == r.n_regions
    g, reg, inner = p.handle(xyz, regions=True)
    assert (reg >= 0).all()
    n_bad = int((reg != reg_ref).sum())
    assert n_bad == 0, f"{n_bad} region index mismatches, first at {np.nonzero(reg != reg_ref)[0][:5]}"
    assert (inner == inner_ref).all()
    assert np.abs(g - g_ref).max() < 1e-12
    assert np.abs(np.linalg.norm(g, axis=1) - 1).max() < 1e-12


def test_stereo_matches_reference(eb, ref):
    rng = np.random.default_rng(5)
    xyz = rng.normal(size=(20000, 3))
    g_ref = ref.PointSourcePanner("0+2+0").handle(xyz, regions=False)
    g = eb.PointSourcePanner("0+2+0").handle(xyz)
    assert g.shape == (20000, 2)
    assert np.abs(g - g_ref).max() < 1e-12


def test_one_hot_at_speakers(eb, ref):
    # configure_full_polar_panner (point_source_panner_tests.cpp:421-519)
    for layout in LAYOUTS:
        info = ref.layout_info(layout)
        pos = info["pos"][~info["is_lfe"]]
        g = eb.PointSourcePanner(layout).handle(cart(pos[:, 0], pos[:, 1]))
        assert np.abs(g - np.eye(len(pos))).max() < 1e-9, layout
    g = eb.PointSourcePanner("0+5+0").handle(cart(np.array([15.0, -15.0]), np.zeros(2)))
    assert np.allclose(g[0, [0, 2]], np.sqrt(0.5)) and np.allclose(g[1, [1, 2]], np.sqrt(0.5))


def test_real_positions_differ_from_nominal(eb, ref):
    # speakers moved inside their allowed ranges: hull from nominal, gains from real positions
    info = ref.layout_info("0+5+0")
    pos = info["pos"].copy()
    pos[4] = [115.0, 10.0, 1.0]
    pos[5] = [-105.0, 5.0, 1.0]
    lay = eb.Layout("0+5+0", real_positions=pos)
    p = eb.PointSourcePanner(lay)
    g = p.handle(cart(np.array([115.0]), np.array([10.0])))
    assert abs(g[0, 3] - 1.0) < 1e-9
    rng = np.random.default_rng(1)
    xyz = rng.normal(size=(1000, 3))
    g = p.handle(xyz)
    assert np.abs(np.linalg.norm(g, axis=1) - 1).max() < 1e-12 and (g >= 0).all()

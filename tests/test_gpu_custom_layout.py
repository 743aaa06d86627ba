"""Layouts outside the BS.2051 table: ear::Layout(name, channels) with a name that has
no tabulated hull -> the panner triangulates the nominal positions itself (ear::convex_hull).
Point-source gains and region indices, and Objects gains, against the compiled reference's golden
vectors (tests/golden/custom.npz) and against the reference live."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KEYS = ["renamed_4+5+0", "ring5_upper4", "mid7_upper2_bottom3"]


def _channels(g, key):
    return dict(names=[str(n) for n in g[f"{key}/names"]], pos=g[f"{key}/pos"], nominal=g[f"{key}/nominal"],
                az_range=g[f"{key}/az_range"], el_range=g[f"{key}/el_range"], is_lfe=g[f"{key}/is_lfe"])


@pytest.mark.parametrize("key", KEYS)
def test_custom_layout_panner_golden(eb, key):
    g = np.load(os.path.join(GOLDEN, "custom.npz"))
    psp = eb.PointSourcePanner(eb.Layout(key, channels=_channels(g, key)))
    gains, reg, inner = psp.handle(g[f"{key}/xyz"], regions=True)
    assert np.array_equal(reg, g[f"{key}/region"])
    assert np.array_equal(inner, g[f"{key}/inner"])
    assert np.abs(gains - g[f"{key}/gains"]).max() < 1e-12


@pytest.mark.parametrize("key", KEYS)
def test_custom_layout_objects_golden(eb, key):
    g = np.load(os.path.join(GOLDEN, "custom.npz"))
    gc = eb.GainCalculatorObjects(eb.Layout(key, channels=_channels(g, key)))
    d, f = gc.calculate(*g[f"{key}/obj_in"])
    assert max(np.abs(d - g[f"{key}/obj_direct"]).max(), np.abs(f - g[f"{key}/obj_diffuse"]).max()) < 1e-5


def test_custom_layout_live(eb, ref):
    # a layout made up here: eleven loudspeakers on three rings, seeded probe directions
    names = ["M+000", "M+060", "M-060", "M+120", "M-120", "M+180", "U+030", "U-030", "U+110", "U-110", "B+000"]
    pos = np.array([[0, 0, 1], [60, 0, 1], [-60, 0, 1], [120, 0, 1], [-120, 0, 1], [180, 0, 1],
                    [30, 32, 1], [-30, 32, 1], [110, 28, 1], [-110, 28, 1], [0, -28, 1]], dtype=np.float64)
    lay_ref = ref.custom_layout("three_rings", names, pos)
    rng = np.random.default_rng(41)
    xyz = rng.normal(size=(50000, 3))
    xyz /= np.linalg.norm(xyz, axis=1, keepdims=True)
    r = ref.PointSourcePanner(lay_ref)
    g_ref, reg_ref, inner_ref = r.handle(xyz)
    psp = eb.PointSourcePanner(eb.Layout("three_rings", channels=dict(names=names, pos=pos)))
    assert psp.num_regions == r.n_regions
    gains, reg, inner = psp.handle(xyz, regions=True)
    assert np.array_equal(reg, reg_ref) and np.array_equal(inner, inner_ref)
    assert np.abs(gains - g_ref).max() < 1e-12
    assert np.abs((gains ** 2).sum(axis=1) - 1.0).max() < 1e-12  # power normalised everywhere


def test_bs2051_name_with_custom_channels_uses_the_table(eb, ref):
    # ear::Layout("4+5+0", channels moved inside their ranges): the tabulated facets, as libear
    info = ref.layout_info("4+5+0")
    pos = info["pos"].copy()
    pos[0, 0], pos[1, 0] = 25.0, -33.0
    lay_ref = ref.custom_layout("4+5+0", info["names"], pos, info["nominal"], info["ranges"][:, :2],
                                info["ranges"][:, 2:], info["is_lfe"])
    rng = np.random.default_rng(43)
    xyz = rng.normal(size=(20000, 3))
    g_ref, reg_ref, _ = ref.PointSourcePanner(lay_ref).handle(xyz)
    lay = eb.Layout("4+5+0", channels=dict(names=info["names"], pos=pos, nominal=info["nominal"],
                                           az_range=info["ranges"][:, :2], el_range=info["ranges"][:, 2:],
                                           is_lfe=info["is_lfe"].astype(np.int32)))
    gains, reg, _ = eb.PointSourcePanner(lay).handle(xyz, regions=True)
    assert np.array_equal(reg, reg_ref) and np.abs(gains - g_ref).max() < 1e-12


def test_degenerate_custom_layout_fails_like_libear(eb):
    # two loudspeakers at the same place: "collinear points in convex hull" (ear::internal_error)
    names = ["M+000", "M+030", "M-030", "M+110", "M-110"]
    pos = [[0, 0, 1], [0, 0, 1], [-30, 0, 1], [110, 0, 1], [-110, 0, 1]]
    with pytest.raises(eb.InternalError):
        eb.PointSourcePanner(eb.Layout("broken", channels=dict(names=names, pos=pos)))

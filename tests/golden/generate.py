#!/usr/bin/env python
"""Generate tests/golden/*.npz from the compiled, UNMODIFIED libear reference
(oracle/_ref/libear_ref.so, built by oracle/Makefile from /root/reference).

Run in the build container (where /root/reference exists):

    make -C oracle && python tests/golden/generate.py

The fixtures are small seeded input/output pairs for every row of the hot path; they pin
the numpy restatement (oracle/ear_oracle.py) on machines where the reference cannot be
built, and the CUDA path on the GPU box.  Known-answer values that the reference's own
tests hold are stored next to them (kat.npz) with the test file:line they come from.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref  # noqa: E402

LAYOUTS = ["0+2+0", "0+5+0", "2+5+0", "4+5+0", "4+5+1", "3+7+0", "4+9+0", "9+10+3", "0+7+0", "4+7+0"]


def cart(az, el, d=1.0):
    az, el = np.radians(az), np.radians(el)
    return np.stack([np.sin(-az) * np.cos(el) * d, np.cos(-az) * np.cos(el) * d, np.sin(el) * d], -1)


def sphere(rng, n):
    v = rng.normal(size=(n, 3))
    return v / np.linalg.norm(v, axis=1, keepdims=True)


def gen_psp():
    """Point-source panner: the 21 x 11 az/el grid of the reference's test_all_layouts
    (tests/point_source_panner_tests.cpp:341-419), every loudspeaker position, and random
    directions, for all ten layouts: gains + region index + inner triplet index."""
    out = {}
    rng = np.random.default_rng(11)
    az, el = np.meshgrid(np.linspace(-180, 180, 21), np.linspace(-90, 90, 11))
    grid = cart(az.ravel(), el.ravel())
    for name in LAYOUTS:
        info = ref.layout_info(name)
        spk = np.array([cart(p[0], p[1]) for p, l in zip(info["pos"], info["is_lfe"]) if not l])
        xyz = np.concatenate([grid, spk, sphere(rng, 120) * rng.uniform(0.1, 2.0, (120, 1))])
        g, reg, inner = ref.PointSourcePanner(name).handle(xyz)
        out[f"{name}/xyz"] = xyz
        out[f"{name}/gains"] = g
        out[f"{name}/region"] = reg
        out[f"{name}/inner"] = inner
    np.savez_compressed(os.path.join(HERE, "psp.npz"), **out)


def gen_extent():
    """PolarExtent::handle on 9+10+3 and 0+5+0 with the distribution of the reference's
    same_as_reference test (tests/extent_tests.cpp:136-166) plus depth/distance: the
    dispatched SIMD core, the scalar core and the double-precision test restatement
    (tests/reference/extent.cpp)."""
    out = {}
    rng = np.random.default_rng(12)
    for name, n in (("9+10+3", 96), ("0+5+0", 32), ("4+7+0", 32)):
        xyz = sphere(rng, n) * rng.uniform(0.2, 1.5, (n, 1))
        w, h = rng.uniform(0, 360, n), rng.uniform(0, 360, n)
        depth = np.where(rng.uniform(size=n) < 0.5, 0.0, rng.uniform(0, 1, n))
        # a few structured cases: zero extent, circular, full sphere, tall, tiny
        w[:6] = [0, 30, 360, 10, 5, 180]
        h[:6] = [0, 30, 360, 200, 0, 180]
        depth[:6] = 0
        xyz[0] /= np.linalg.norm(xyz[0])
        out[f"{name}/xyz"], out[f"{name}/width"], out[f"{name}/height"], out[f"{name}/depth"] = xyz, w, h, depth
        out[f"{name}/simd"] = ref.Extent(name, 0).handle(xyz, w, h, depth)
        out[f"{name}/scalar"] = ref.Extent(name, 1).handle(xyz, w, h, depth)
        unit = xyz / np.linalg.norm(xyz, axis=1, keepdims=True)
        out[f"{name}/double_nodepth_unit"] = ref.Extent(name, 2).handle(unit, w, h, 0.0)
        out[f"{name}/simd_nodepth_unit"] = ref.Extent(name, 0).handle(unit, w, h, 0.0)
    out["isa"] = np.array(ref.extent_isa())
    np.savez_compressed(os.path.join(HERE, "extent.npz"), **out)


def gen_objects():
    """GainCalculatorObjects::calculate: BASELINE cfg 1 distribution (point sources on 0+5+0),
    cfg 2 distribution (extent objects on 9+10+3) and 4+7+0 (cfg 5), float and double."""
    out = {}
    rng = np.random.default_rng(13)
    for name, n, extent in (("0+5+0", 128, False), ("9+10+3", 96, True), ("4+7+0", 48, True), ("0+2+0", 24, True)):
        az = rng.uniform(-180, 180, n)
        el = np.degrees(np.arcsin(rng.uniform(-1, 1, n)))
        if extent:
            dist = rng.uniform(0.2, 1.5, n)
            w, h = rng.uniform(0, 360, n), rng.uniform(0, 360, n)
            depth = np.where(rng.uniform(size=n) < 0.5, 0.0, rng.uniform(0, 1, n))
        else:
            dist, w, h, depth = np.ones(n), np.zeros(n), np.zeros(n), np.zeros(n)
        gain, diffuse = rng.uniform(0, 1, n), rng.uniform(0, 1, n)
        calc = ref.Objects(name)
        d64, f64 = calc.calculate(az, el, dist, w, h, depth, gain, diffuse, dtype=np.float64)
        d32, f32 = calc.calculate(az, el, dist, w, h, depth, gain, diffuse, dtype=np.float32)
        out[f"{name}/in"] = np.stack([az, el, dist, w, h, depth, gain, diffuse])
        out[f"{name}/direct64"], out[f"{name}/diffuse64"] = d64, f64
        out[f"{name}/direct32"], out[f"{name}/diffuse32"] = d32, f32
        out[f"{name}/region"] = ref.PointSourcePanner(name).handle(cart(az, el, dist))[1]
    np.savez_compressed(os.path.join(HERE, "objects.npz"), **out)


def gen_interp():
    """GainInterpolator<LinearInterpMatrix/Vector/Single>::process: ramps, steps (equal
    sample indices), constant head/tail, equal neighbouring points, call splitting and a
    non-zero block_start (tests/gain_interpolator_tests.cpp:104-257)."""
    out = {}
    rng = np.random.default_rng(14)
    N, M, T = 6, 5, 1500
    x = rng.normal(0, 0.25, (N, T)).astype(np.float32)
    samples = np.array([100, 300, 300, 612, 800, 1124, 1124, 1400], dtype=np.int64)
    g = rng.uniform(0, 1, (len(samples), N, M)).astype(np.float32)
    g[5] = g[4]  # constant between 800 and 1124
    out["matrix/x"], out["matrix/samples"], out["matrix/gains"] = x, samples, g
    out["matrix/out"] = ref.interp_matrix(x, samples, g)
    out["matrix/out_calls_of_75"] = ref.interp_matrix(x, samples, g, call_size=75)
    out["matrix/out_start_250"] = ref.interp_matrix(x[:, :1000], samples, g, block_start0=250)
    out["vector/out"] = ref.interp_vector(x[0], samples, g[:, 0, :])
    out["single/out"] = ref.interp_single(x[0], samples, g[:, 0, 0])
    # BASELINE cfg 3 shape in miniature: 128 -> 14, ramps every 512 samples
    # (input stored as float16 -- exactly representable -- to keep the fixture small)
    N, M, T = 128, 14, 2048
    x16 = np.clip(rng.normal(0, 0.25, (N, T)), -1, 1).astype(np.float16)
    samples = np.arange(0, T + 512, 512, dtype=np.int64)
    g = (rng.uniform(0, 1, (len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
    out["cfg3/x16"], out["cfg3/samples"], out["cfg3/gains"] = x16, samples, g
    out["cfg3/out"] = ref.interp_matrix(x16.astype(np.float32), samples, g)
    np.savez_compressed(os.path.join(HERE, "interp.npz"), **out)
    # the same shape with float32 noise (NOT float16-exact: x_lo != 0, so both cross terms of a
    # split-TF32 contraction are exercised), libear's output and the float64 result of the same ramp
    N, M, T = 128, 14, 1024
    x = np.clip(rng.normal(0, 0.25, (N, T)), -1, 1).astype(np.float32)
    samples = np.arange(0, T + 512, 512, dtype=np.int64)
    g = (rng.uniform(0, 1, (len(samples), N, M)) / np.sqrt(N)).astype(np.float32)
    t = np.arange(T)
    k0 = np.searchsorted(samples, t, side="right") - 1
    p_ = (t - samples[k0]) / 512.0
    gg = g[k0].astype(np.float64) * (1 - p_)[:, None, None] + g[k0 + 1].astype(np.float64) * p_[:, None, None]
    np.savez_compressed(os.path.join(HERE, "interp_f32.npz"), x=x, samples=samples, gains=g,
                        libear=ref.interp_matrix(x, samples, g),
                        float64=np.einsum("nt,tnm->mt", x.astype(np.float64), gg))


def gen_convolver():
    """BlockConvolver::process scenarios of tests/block_convolver_tests.cpp:208-356:
    multi-partition filters, crossfades, fades to/from silence, zero blocks."""
    out = {}
    rng = np.random.default_rng(15)
    B, nb, blocks = 64, 4, 14
    x = rng.normal(0, 0.25, B * blocks).astype(np.float32)
    x[5 * B:6 * B] = 0.0  # an all-zero input block
    f0 = rng.normal(0, 0.3, 200).astype(np.float32)   # 4 partitions
    f1 = rng.normal(0, 0.3, 100).astype(np.float32)   # 2 partitions
    out["B"], out["nb"], out["x"], out["f0"], out["f1"] = np.array(B), np.array(nb), x, f0, f1
    c = ref.Convolver(B, nb)
    a, _ = c.add_filter(f0)
    b, _ = c.add_filter(f1)
    c.set_filter(a)
    y = [c.process(x[:3 * B])]
    c.crossfade_filter(b)
    y.append(c.process(x[3 * B:7 * B]))
    c.crossfade_filter(None)                      # fade down to silence
    y.append(c.process(x[7 * B:9 * B]))
    y.append(c.process(None, 2))                  # NULL input blocks
    c.crossfade_filter(a)                         # fade up from silence
    y.append(c.process(x[11 * B:]))
    out["y"] = np.concatenate(y)
    out["script"] = np.array("set f0; 3 blocks; xfade f1; 4 blocks; xfade none; 2 blocks; "
                             "2 null blocks; xfade f0; 3 blocks")
    # BASELINE cfg 4 in miniature: decorrelator filters on 9+10+3, B = 512, one partition
    dec = ref.decorrelators("9+10+3", np.float32)
    xs = rng.normal(0, 0.25, (3, 512 * 6)).astype(np.float32)
    ys = []
    for ch, filt in zip(xs, dec[[0, 7, 23]]):
        c = ref.Convolver(512, 1)
        fid, _ = c.add_filter(filt)
        c.set_filter(fid)
        ys.append(c.process(ch))
    out["cfg4/x"], out["cfg4/channels"], out["cfg4/y"] = xs, np.array([0, 7, 23]), np.array(ys)
    np.savez_compressed(os.path.join(HERE, "convolver.npz"), **out)


def gen_decorrelators():
    out = {}
    for name in ("0+5+0", "4+7+0", "9+10+3"):
        out[f"{name}/f32"] = ref.decorrelators(name, np.float32)
        out[f"{name}/f64"] = ref.decorrelators(name, np.float64)
    out["delay"] = np.array(ref.decorrelator_delay())
    np.savez_compressed(os.path.join(HERE, "decorrelators.npz"), **out)


def gen_hoa():
    """GainCalculatorHOA::calculate: ACN 0..15 on 4+7+0 (cfg 5) and 9+10+3, all three
    normalisations; first/last t-design point (tests/hoa_tests.cpp:7-16)."""
    out = {}
    orders = [n for n in range(4) for _ in range(2 * n + 1)]
    degrees = [m for n in range(4) for m in range(-n, n + 1)]
    out["orders"], out["degrees"] = np.array(orders), np.array(degrees)
    for name in ("4+7+0", "9+10+3", "0+5+0"):
        for norm in ("SN3D", "N3D", "FuMa"):
            out[f"{name}/{norm}"] = ref.hoa_calc(name, orders, degrees, norm)[0]
    pts = ref.hoa_points()
    out["points_first_last"] = pts[[0, -1]]
    out["n_points"] = np.array(len(pts))
    np.savez_compressed(os.path.join(HERE, "hoa.npz"), **out)


def direct_speakers_cases():
    """The reference's own test cases (tests/gain_calculator_direct_speakers_tests.cpp:31-339:
    label / URN matching, LFE by name or frequency, bounds, point-source fallback, mapping rules
    with common-definitions pack ids, the not_implemented / adm_error throws) plus a seeded random
    sweep over all layouts."""
    cases = [
        dict(layout="4+7+0", labels=["M+000"]),
        dict(layout="4+7+0", labels=["urn:itu:bs:2051:0:speaker:M+030"], az=30),
        dict(layout="4+7+0", labels=["LFE1"], el=-30),
        dict(layout="4+7+0", labels=["foo"], el=-30, low_pass=120.0),          # LFE by frequency
        dict(layout="4+7+0", labels=["foo"], el=-30, low_pass=250.0),          # not an LFE frequency
        dict(layout="4+7+0", labels=["LFE1"], el=-30, low_pass=120.0, high_pass=20.0),
        dict(layout="4+7+0", labels=["foo"], az=12, el=5),                      # point-source fallback
        dict(layout="4+7+0", labels=["foo"], az=30, bounds={"azMin": 25, "azMax": 35, "elMin": -5, "elMax": 5}),
        dict(layout="4+7+0", labels=["foo"], az=70, el=28,
             bounds={"azMin": 40, "azMax": 100, "elMin": 20, "elMax": 40}),
        dict(layout="4+7+0", labels=["foo"], az=0, bounds={"azMin": -180, "azMax": 180}),   # ambiguous -> panned
        dict(layout="4+7+0", labels=[], az=-110),
        dict(layout="4+7+0", labels=[], az=180, el=90, bounds={"elMin": 80, "elMax": 90}),
        dict(layout="9+10+3", labels=["U+045"], az=45, el=30),
        dict(layout="9+10+3", labels=["B-045"], az=-45, el=-30),
        dict(layout="9+10+3", labels=["LFER"], az=-45, el=-30),
        dict(layout="9+10+3", labels=["LFE"], az=45, el=-30),
        dict(layout="9+10+3", labels=["foo"], el=-30, low_pass=100.0, az=-50),
        dict(layout="0+5+0", labels=["M+110"], az=110),
        dict(layout="0+5+0", labels=["LFE2"], az=-45, el=-30),                   # no LFE2: LFE rule -> LFE1
        dict(layout="0+2+0", labels=["LFE1"], az=45, el=-30),                    # no LFE at all: silence
        dict(layout="0+5+0", labels=["U+030"], az=30, el=30, pack="AP_00010005"),  # 4+5+0 -> 0+5+0 rule
        dict(layout="0+5+0", labels=["M+090"], az=90, pack="AP_00010008"),
        dict(layout="0+5+0", labels=["urn:itu:bs:2051:1:speaker:M+135"], az=135, pack="AP_00010017"),
        dict(layout="0+2+0", labels=["M+000"], pack="AP_00010003"),
        dict(layout="0+2+0", labels=["M+110"], az=110, pack="AP_00010003"),
        dict(layout="4+5+0", labels=["UH+180"], az=180, el=45, pack="AP_00010009"),
        dict(layout="4+5+0", labels=["LFE2"], az=-45, el=-30, pack="AP_00010009"),
        dict(layout="4+7+0", labels=["T+000"], el=90, pack="AP_00010009"),
        dict(layout="0+5+0", labels=["U+030"], az=30, el=30, pack="AP_ffffffff"),   # unknown pack
        dict(layout="0+5+0", labels=["U+030"], az=30, el=30, pack="urn:itu:bs:2051:0:pack:4+5+0"),
        dict(layout="4+7+0", labels=["left"], az=30, substitutions={"left": "M+030", "LFE": "LFE2"}),
        dict(layout="4+7+0", labels=["LFE"], el=-30, substitutions={"left": "M+030", "LFE": "LFE2"}),
        dict(layout="4+7+0", labels=[], pack="AP_00010003"),                      # adm_error
        dict(layout="4+7+0", labels=["foo"], cartesian=True),                     # not_implemented
        dict(layout="4+7+0", labels=["foo"], az=5, screen_edge_lock=1),           # not_implemented
        dict(layout="4+7+0", labels=["M+000"], screen_edge_lock=2),               # label match wins first
    ]
    rng = np.random.default_rng(17)
    names = ref.layout_names()
    pool = sorted({n for L in names for n in ref.layout_info(L)["names"]}) + [
        "LFE", "LFEL", "LFER", "foo", "urn:itu:bs:2051:0:speaker:M+030", "urn:itu:bs:2051:12:speaker:U-110",
        "urn:itu:bs:2051::speaker:M+030", "urn:itu:bs:2051:0:speaker:"]
    packs = [None, None, "AP_00010001", "AP_00010002", "AP_00010003", "AP_00010004", "AP_00010005",
             "AP_00010007", "AP_00010008", "AP_00010009", "AP_0001000c", "AP_0001000f", "AP_00010010",
             "AP_00010017"]
    for _ in range(400):
        c = dict(layout=names[rng.integers(len(names))],
                 labels=[pool[rng.integers(len(pool))] for _ in range(rng.integers(0, 3))],
                 az=float(rng.choice([0, 30, -30, 110, -110, 45, 90, 180, -180, rng.uniform(-180, 180)])),
                 el=float(rng.choice([0, 0, 30, -30, 90, rng.uniform(-90, 90)])),
                 dist=float(rng.choice([1.0, 1.0, 0.5])))
        if rng.uniform() < 0.4:
            b = {}
            if rng.uniform() < 0.7:
                b["azMin"], b["azMax"] = c["az"] - float(rng.uniform(0, 40)), c["az"] + float(rng.uniform(0, 40))
            if rng.uniform() < 0.7:
                b["elMin"], b["elMax"] = c["el"] - float(rng.uniform(0, 20)), c["el"] + float(rng.uniform(0, 20))
            if rng.uniform() < 0.3:
                b["distMin"], b["distMax"] = 0.0, 2.0
            c["bounds"] = b
        if rng.uniform() < 0.2:
            c["low_pass"] = float(rng.choice([80.0, 120.0, 200.0, 500.0]))
        if rng.uniform() < 0.05:
            c["high_pass"] = 40.0
        pk = packs[rng.integers(len(packs))]
        if pk and c["labels"]:
            c["pack"] = pk
        cases.append(c)
    return cases


def gen_direct_speakers():
    import json
    cases = direct_speakers_cases()
    out = {"n": np.array(len(cases))}
    for i, c in enumerate(cases):
        try:
            g, codes = ref.direct_speakers_calc(c["layout"], c.get("labels", []), c.get("az", 0.0), c.get("el", 0.0),
                                                c.get("dist", 1.0), c.get("bounds"), c.get("low_pass"),
                                                c.get("high_pass"), c.get("pack"), c.get("substitutions"),
                                                c.get("cartesian", False), c.get("screen_edge_lock", 0), codes=True)
            err = 0
        except ref.RefError as e:
            g, codes, err = np.zeros(0), [], e.code
        out[f"{i}/case"] = np.array(json.dumps(c))
        out[f"{i}/gains"] = g
        out[f"{i}/warnings"] = np.array(codes, dtype=np.int32)
        out[f"{i}/error"] = np.array(err)
    np.savez_compressed(os.path.join(HERE, "direct_speakers.npz"), **out)


def gen_delay():
    rng = np.random.default_rng(16)
    x = rng.normal(0, 0.25, (3, 1000)).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "delay.npz"), x=x, delay=np.array(255),
                        y=ref.delay_process(x, 255, call_size=128))


def conversion_inputs():
    """Seeded positions / extents plus the cases of the reference's own conversion tests
    (tests/conversion_tests.cpp:52-160): the az x el x dist loop, poles, centre, sector corners."""
    rng = np.random.default_rng(17)
    n = 400
    polar = [[az, el, d] for az in (0.0, -10.0, 10.0, 90.0, -90.0, 150.0, -150.0)
             for el in (0.0, -10.0, 10.0, -45.0, 45.0, -90.0, 90.0) for d in (0.5, 1.0)]
    polar += [[0.0, s * 90.0, d] for s in (-1.0, 1.0) for d in (0.5, 1.0, 2.0)]
    polar += [[az, el, 0.0] for az in (-90.0, 0.0, 90.0) for el in (-90.0, 0.0, 90.0)]
    polar += [[az, el, 1.0] for az in (0.0, -30.0, -110.0, 110.0, 30.0, 180.0, -180.0, 360.0, -400.0)
              for el in (0.0, 30.0, -30.0, 30.000001)]
    polar += [[10.0, 20.0, 0.3]]
    polar = np.array(polar)
    polar = np.concatenate([polar, np.stack([rng.uniform(-180, 180, n), rng.uniform(-90, 90, n),
                                             rng.uniform(0, 2, n)], 1)])
    polar_ext = np.stack([rng.uniform(0, 360, len(polar)), rng.uniform(0, 360, len(polar)),
                          rng.uniform(0, 1, len(polar))], 1)
    polar_ext[:20] = 0.0
    polar_ext[-1 - n] = [40.0, 50.0, 0.6]  # conversion_reference, tests/conversion_tests.cpp:163
    cartp = [[0.0, 0.0, s * d] for s in (-1.0, 1.0) for d in (0.5, 1.0, 2.0)] + [[0.0, 0.0, 0.0]]
    cartp += [[x, y, z] for x in (-1.0, 0.0, 1.0) for y in (-1.0, 0.0, 1.0) for z in (-1.0, 0.0, 0.5, 1.0)
              if (x, y) != (0.0, 0.0)]
    cartp += [[1e-11, 1e-11, 0.5], [1e-9, 0.0, 0.5], [0.9, 0.8, 0.1]]
    cartp = np.array(cartp)
    cartp = np.concatenate([cartp, rng.uniform(-1.2, 1.2, (n, 3))])
    cart_ext = rng.uniform(0, 1, (len(cartp), 3))
    cart_ext[:10] = 0.0
    cart_ext[-1 - n] = [0.3, 0.5, 0.4]  # conversion_reference, tests/conversion_tests.cpp:175
    return polar, polar_ext, cartp, cart_ext


def gen_conversion():
    polar, polar_ext, cartp, cart_ext = conversion_inputs()
    p2c_pos, p2c_ext = ref.conversion(polar, polar_ext, to_polar=False)
    c2p_pos, c2p_ext = ref.conversion(cartp, cart_ext, to_polar=True)
    np.savez_compressed(os.path.join(HERE, "conversion.npz"), polar=polar, polar_ext=polar_ext, p2c_pos=p2c_pos,
                        p2c_ext=p2c_ext, cart=cartp, cart_ext=cart_ext, c2p_pos=c2p_pos, c2p_ext=c2p_ext)


def custom_layouts():
    """Layouts whose names are not in BS.2051 (ear::Layout(name, channels)): the panner builds
    their hull itself (convex_hull).  Channel names still follow BS.2051 (the extra-speaker rule
    reads the layers from them); LFE channels are dropped by every calculator."""
    out = {}
    # 4+5+0's channels under another name, real positions moved inside the allowed ranges
    info = ref.layout_info("4+5+0")
    pos = info["pos"].copy()
    pos[0, 0], pos[1, 0], pos[4, 0], pos[6, 1] = 25.0, -35.0, 118.0, 38.0
    out["renamed_4+5+0"] = ref.custom_layout("renamed_4+5+0", info["names"], pos, info["nominal"],
                                             info["ranges"][:, :2], info["ranges"][:, 2:], info["is_lfe"])
    # mid ring of five at non-standard azimuths + four upper speakers; nominal = real
    names = ["M+000", "M+045", "M-045", "M+135", "M-135", "U+045", "U-045", "U+135", "U-135"]
    pos = [[0, 0, 1], [45, 0, 1], [-45, 0, 1], [135, 0, 1], [-135, 0, 1],
           [45, 35, 1], [-45, 35, 1], [135, 35, 1], [-135, 35, 1]]
    out["ring5_upper4"] = ref.custom_layout("ring5_upper4", names, pos)
    # seven mid, two upper (extra upper speakers appear at the back), three bottom, one LFE
    names = ["M+000", "M+030", "M-030", "M+090", "M-090", "M+135", "M-135", "U+045", "U-045",
             "B+000", "B+045", "B-045", "LFE1"]
    pos = [[0, 0, 1], [30, 0, 1], [-30, 0, 1], [90, 0, 1], [-90, 0, 1], [135, 0, 1], [-135, 0, 1],
           [45, 30, 1], [-45, 30, 1], [0, -30, 1], [45, -30, 1], [-45, -30, 1], [45, -30, 1]]
    lfe = [0] * 12 + [1]
    out["mid7_upper2_bottom3"] = ref.custom_layout("mid7_upper2_bottom3", names, pos, is_lfe=lfe)
    return out


def gen_custom():
    """Custom-layout path: ear::convex_hull facets for a set of point clouds, and
    point-source gains / region indices / Objects gains on layouts with non-BS.2051 names."""
    out = {}
    rng = np.random.default_rng(18)
    clouds = []
    for name in LAYOUTS[1:]:
        info = ref.layout_info(name)
        nom = info["nominal"][~info["is_lfe"]]
        pts = [cart(nom[:, 0], nom[:, 1]), [[0.0, 0.0, -1.0]]]
        if "T+000" not in info["names"] and "UH+180" not in info["names"]:
            pts.append([[0.0, 0.0, 1.0]])
        clouds.append(np.concatenate(pts))
    clouds += [sphere(rng, n) for n in (4, 5, 6, 9, 14, 20)]
    clouds.append(cart(np.array([45.0, -45, 135, -135, 45, -45, 135, -135]),
                       np.array([35.0, 35, 35, 35, -35, -35, -35, -35])))  # a box: six quads
    out["n_clouds"] = np.array(len(clouds))
    for k, pts in enumerate(clouds):
        facets = ref.convex_hull(pts)
        out[f"hull{k}/xyz"] = pts
        out[f"hull{k}/facets"] = np.array([x for f in facets for x in [len(f)] + f], dtype=np.int32)
    az, el = np.meshgrid(np.linspace(-180, 180, 21), np.linspace(-90, 90, 11))
    grid = cart(az.ravel(), el.ravel())
    lays = custom_layouts()
    out["layouts"] = np.array(sorted(lays))
    for key, lay in lays.items():
        for f in ("pos", "nominal", "az_range", "el_range", "is_lfe"):
            out[f"{key}/{f}"] = lay[f]
        out[f"{key}/names"] = np.array(lay["names"])
        keep = lay["is_lfe"] == 0
        spk = cart(lay["pos"][keep, 0], lay["pos"][keep, 1])
        xyz = np.concatenate([grid, spk, sphere(rng, 300) * rng.uniform(0.1, 2.0, (300, 1))])
        g, reg, inner = ref.PointSourcePanner(lay).handle(xyz)
        out[f"{key}/xyz"], out[f"{key}/gains"], out[f"{key}/region"], out[f"{key}/inner"] = xyz, g, reg, inner
        n = 48
        obj = np.stack([rng.uniform(-180, 180, n), np.degrees(np.arcsin(rng.uniform(-1, 1, n))),
                        rng.uniform(0.2, 1.5, n), rng.uniform(0, 360, n), rng.uniform(0, 360, n),
                        np.where(rng.uniform(size=n) < 0.5, 0.0, rng.uniform(0, 1, n)),
                        rng.uniform(0, 1, n), rng.uniform(0, 1, n)])
        obj[3:6, :8] = 0.0  # some point sources
        d, f = ref.Objects(lay).calculate(*obj, dtype=np.float32)
        out[f"{key}/obj_in"], out[f"{key}/obj_direct"], out[f"{key}/obj_diffuse"] = obj, d, f
    np.savez_compressed(os.path.join(HERE, "custom.npz"), **out)


def gen_kat():
    """Known answers held by the reference's own tests (not produced by running anything)."""
    kat = {
        # tests/decorrelate_tests.cpp:19-28 -- designDecorrelatorBasic(7, 512) at these taps
        "decorrelator7_taps": np.array([0, 1, 255, 256, 510, 511]),
        "decorrelator7_values": np.array([-0.1124280906086625, -0.00944671630601479, 0.057714955000898516,
                                          -0.018996037984052125, 0.08336121588594464,
                                          -0.012216595581941523]),
        # tests/point_source_panner_tests.cpp:421-519 -- 0+5+0, +-15 degrees: (sqrt 1/2, sqrt 1/2)
        "psp_0+5+0_az15": np.array([np.sqrt(0.5), 0.0, np.sqrt(0.5), 0.0, 0.0]),
        # tests/hoa_tests.cpp:7-16 -- first / last point of the 5200-point design
        "hoa_point_first": np.array([-0.007238307489963788, 0.8575241661297168, 0.5143927598714958]),
        "hoa_point_last": np.array([-0.5784605915309438, 0.5246941913018843, -0.6245633271812653]),
        "hoa_n_points": np.array(5200),
        # tests/conversion_tests.cpp:162-185 -- extentPolarToCart({10, 20, 0.3}, {40, 50, 0.6}) and
        # extentCartToPolar({0.9, 0.8, 0.1}, {0.3, 0.5, 0.4}): position then extent
        "conversion_p2c": np.array([-0.08972503721988338, 0.3, 0.1732050807568877, 0.35166171614357594,
                                    0.4470181645863707, 0.5762749096794243]),
        "conversion_c2p": np.array([-34.85107611658391, 4.226794497273273, 0.9000000000000001,
                                    76.50724453298275, 104.9708107421662, 0.1756348204517474]),
    }
    np.savez_compressed(os.path.join(HERE, "kat.npz"), **kat)


if __name__ == "__main__":
    if not ref.available():
        raise SystemExit("build oracle/_ref first: make -C oracle")
    for fn in (gen_psp, gen_extent, gen_objects, gen_interp, gen_convolver, gen_decorrelators, gen_hoa,
               gen_direct_speakers, gen_delay, gen_conversion, gen_custom, gen_kat):
        fn()
        print("wrote", fn.__name__[4:])

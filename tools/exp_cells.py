"""Experiment: how many (pass, cell) pairs of the extent grid are 'mixed' (neither all weights
exactly 1 nor all exactly 0) when the 1,652 virtual sources are clustered into compact cells
and a cell is classified from the distance of its centre to the stadium spine +- its radius.
Pure numpy, cfg-2 metadata distribution.  Usage: python tools/exp_cells.py [cell]"""
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle import ear_oracle as eo  # noqa: E402

PI = np.pi


def cluster(points, cell, rows_per_band):
    """bands of grid rows, points of a band sorted by azimuth, cut into cells"""
    el = np.degrees(np.arcsin(np.clip(points[:, 2], -1, 1)))
    az = np.degrees(np.arctan2(points[:, 0], points[:, 1]))
    rows = np.round((el + 90) / 5).astype(int)
    order = []
    r = 0
    nrows = rows.max() + 1
    bands = []
    # poles: merge rows until a band has at least `cell` points
    while r < nrows:
        r1 = min(nrows, r + rows_per_band)
        idx = np.where((rows >= r) & (rows < r1))[0]
        bands.append(idx)
        r = r1
    for idx in bands:
        order.extend(idx[np.argsort(az[idx], kind="stable")])
    order = np.array(order)
    cells = [order[i:i + cell] for i in range(0, len(order), cell)]
    return cells


def cell_caps(points, cells):
    cs, rs = [], []
    for c in cells:
        p = points[c]
        ctr = p.mean(axis=0)
        ctr /= np.linalg.norm(ctr)
        # a few rounds of "move towards the farthest point" (approximate min enclosing cap)
        for it in range(200):
            d = p @ ctr
            j = np.argmin(d)
            ctr = ctr + 0.02 * (p[j] - ctr)
            ctr /= np.linalg.norm(ctr)
        cs.append(ctr)
        rs.append(np.arccos(np.clip((p @ ctr).min(), -1, 1)))
    return np.array(cs), np.array(rs)


def pass_params(rng, n):
    out = []
    for _ in range(n):
        v = rng.normal(size=3)
        v /= np.linalg.norm(v)
        dist = rng.uniform(0.2, 1.5)
        w, h = rng.uniform(0, 360, 2)
        depth = 0.0 if rng.uniform() < 0.5 else rng.uniform(0, 1)
        ds = [dist] if depth == 0 else [max(dist - depth / 2, 0), max(dist + depth / 2, 0)]
        for d in ds:
            out.append((v * dist, eo.extent_mod(w, d), eo.extent_mod(h, d)))
    return out


def spine_distance(points, position, width, height):
    """distance (rad) of every point to the stadium spine + the start angle (half height)"""
    width, height = np.radians(max(width, 5.0)) / 2, np.radians(max(height, 5.0)) / 2
    m = eo.calc_basis(position)
    if height > width:
        width, height = height, width
        m = np.array([[0.0, 0.0, 1.0], [0.0, 1.0, 0.0], [-1.0, 0.0, 0.0]]) @ m
    width_full = PI + height
    width_mod = eo.interp(width, [0.0, PI / 2, PI], [0.0, PI / 2, width_full])
    width = eo.interp(height, [0, PI / 4, PI / 2, PI], [width_mod, width_mod, width, width])
    cp = width - height
    t = points @ m.T
    ax = np.abs(t[:, 0])
    straight = (-ax * np.cos(cp) + t[:, 1] * np.sin(cp)) >= 0
    if cp < 1e-6:
        straight[:] = False
    d = np.where(straight, np.arcsin(np.clip(np.abs(t[:, 2]), 0, 1)),
                 np.arccos(np.clip(ax * np.sin(cp) + t[:, 1] * np.cos(cp), -1, 1)))
    return d, height


def main():
    cell = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    rows_per_band = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    pts = eo.extent_grid()
    cells = cluster(pts, cell, rows_per_band)
    ctr, rad = cell_caps(pts, cells)
    print(f"{len(cells)} cells of {cell}; radius deg: mean {np.degrees(rad).mean():.2f} "
          f"max {np.degrees(rad).max():.2f} min {np.degrees(rad).min():.2f}")
    rng = np.random.default_rng(2)
    ps = pass_params(rng, 1500)
    fade = np.radians(10.0)
    mixed_cells, mixed_true, fade_pts, tot = 0, 0, 0, 0
    per_pass = []
    for pos, w, h in ps:
        if max(w, h) <= 1e-9:
            continue
        d, start = spine_distance(pts, pos, w, h)
        dc, _ = spine_distance(ctr, pos, w, h)
        wts = np.clip(1 - (d - start) / fade, 0, 1)
        ones = dc + rad <= start
        zeros = dc - rad >= start + fade
        mixed = ~(ones | zeros)
        # check conservativeness
        for ci in np.where(ones)[0]:
            assert np.all(wts[cells[ci]] >= 1 - 1e-9), (wts[cells[ci]], np.degrees(dc[ci]), np.degrees(rad[ci]))
        for ci in np.where(zeros)[0]:
            assert np.all(wts[cells[ci]] <= 1e-9)
        true_mixed = sum(1 for c in cells if not (np.all(wts[c] >= 1) or np.all(wts[c] <= 0)))
        mixed_cells += mixed.sum()
        mixed_true += true_mixed
        fade_pts += ((wts > 0) & (wts < 1)).sum()
        tot += 1
        per_pass.append(mixed.sum())
    per_pass = np.array(per_pass)
    print(f"passes {tot}: mixed by cap test {mixed_cells / tot / len(cells):.3f}, truly mixed "
          f"{mixed_true / tot / len(cells):.3f}, points on the fade {fade_pts / tot / len(pts):.3f}")
    print("mixed cells per pass: mean %.1f p10 %d p50 %d p90 %d max %d" % (
        per_pass.mean(), *np.percentile(per_pass, [10, 50, 90]).astype(int), per_pass.max()))


if __name__ == "__main__":
    main()

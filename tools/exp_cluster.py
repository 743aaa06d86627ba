"""Experiment: cap radii and the mixed fraction of (pass, cell) pairs for the clustering the
product builds (recursive principal-axis bisection, extent_cells_build) and for refinements of
it.  Pure numpy, cfg-2 metadata distribution.  Usage: python tools/exp_cluster.py [n_passes]"""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tools")
from oracle import ear_oracle as eo  # noqa: E402
import exp_cells as ec  # noqa: E402

CELL = 16


def bisect(pts, idx, cells):
    n = len(idx)
    if n <= CELL:
        cells.append(list(idx))
        return
    p = pts[idx]
    mean = p.mean(axis=0)
    d = p - mean
    cov = d.T @ d
    ax = np.array([0.5773502691896258, 0.5773502691896258 * 1.1, 0.5773502691896258 * 0.9])
    for _ in range(200):
        t = cov @ ax
        nrm = np.linalg.norm(t)
        if nrm < 1e-300:
            break
        ax = t / nrm
    proj = d @ ax
    order = np.argsort(proj, kind="stable")
    n_cells = (n + CELL - 1) // CELL
    left = (n_cells // 2) * CELL
    idx = np.asarray(idx)
    bisect(pts, idx[order[:left]], cells)
    bisect(pts, idx[order[left:]], cells)


def cap(p):
    """bounding cap of the points p: centre, radius (the product's walk towards the farthest point)"""
    ctr = p.sum(axis=0)
    ctr /= np.linalg.norm(ctr)
    best_c, best_r = ctr, np.arccos(np.clip((p @ ctr).min(), -1, 1))
    for it in range(400):
        j = np.argmin(p @ ctr)
        ctr = ctr + (0.5 / (1.0 + 0.05 * it)) * (p[j] - ctr) * 0.1
        ctr /= np.linalg.norm(ctr)
        r = np.arccos(np.clip((p @ ctr).min(), -1, 1))
        if r < best_r:
            best_c, best_r = ctr.copy(), r
    return best_c, best_r


def caps(pts, cells):
    cr = [cap(pts[c]) for c in cells]
    return np.array([c for c, _ in cr]), np.array([r for _, r in cr])


def evaluate(pts, cells, passes, label):
    ctr, rad = caps(pts, cells)
    fade = np.radians(10.0)
    mixed = 0
    for pos, w, h in passes:
        dc, start = ec.spine_distance(ctr, pos, w, h)
        ones = dc + rad <= start
        zeros = dc - rad >= start + fade
        mixed += (~(ones | zeros)).sum()
    deg = np.degrees(rad)
    print(f"{label}: {len(cells)} cells, radius mean {deg.mean():.2f} max {deg.max():.2f} min {deg.min():.2f} | "
          f"mixed cells per pass {mixed / len(passes):.2f} ({mixed / len(passes) / len(cells):.3f})")
    return mixed / len(passes)


def refine(pts, cells, rounds=6):
    """pairs of neighbouring cells re-split along the principal axis of their union and by moving
    single points, whenever that lowers the sum of the two radii"""
    cells = [list(c) for c in cells]
    ctr, rad = caps(pts, cells)
    for rd in range(rounds):
        improved = 0
        for a in range(len(cells)):
            # the six nearest cells
            near = np.argsort(-(ctr @ ctr[a]))[1:7]
            for b in near:
                if b == a:
                    continue
                pool = np.array(cells[a] + cells[b])
                na = len(cells[a])
                best = rad[a] + rad[b]
                best_split = None
                p = pts[pool]
                d = p - p.mean(axis=0)
                # candidate axes: principal axis, centre difference, and a few rotations between them
                _, _, vt = np.linalg.svd(d, full_matrices=False)
                cands = [vt[0], ctr[b] - ctr[a], vt[0] + 0.5 * vt[1], vt[0] - 0.5 * vt[1]]
                for axv in cands:
                    order = np.argsort(d @ axv, kind="stable")
                    for cut in (na, len(pool) - na):
                        if cut <= 0 or cut >= len(pool) or cut > CELL or len(pool) - cut > CELL:
                            continue
                        s1, s2 = pool[order[:cut]], pool[order[cut:]]
                        (c1, r1), (c2, r2) = cap(pts[s1]), cap(pts[s2])
                        if r1 + r2 < best - 1e-9:
                            best, best_split = r1 + r2, (s1, s2, c1, r1, c2, r2)
                if best_split:
                    s1, s2, c1, r1, c2, r2 = best_split
                    cells[a], cells[b] = list(s1), list(s2)
                    ctr[a], rad[a], ctr[b], rad[b] = c1, r1, c2, r2
                    improved += 1
        print(f"  refine round {rd}: {improved} re-splits, mean radius {np.degrees(rad).mean():.2f}")
        if not improved:
            break
    return cells


def main():
    n_passes = int(sys.argv[1]) if len(sys.argv) > 1 else 600
    pts = eo.extent_grid()
    rng = np.random.default_rng(2)
    passes = [p for p in ec.pass_params(rng, n_passes) if max(p[1], p[2]) > 1e-9]
    cells = []
    bisect(pts, np.arange(len(pts)), cells)
    evaluate(pts, cells, passes, "bisection (product)")
    ref = refine(pts, cells)
    evaluate(pts, ref, passes, "refined")


if __name__ == "__main__":
    main()

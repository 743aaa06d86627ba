#!/usr/bin/env python
"""Generates libear_b200/csrc/extent_cells_1652.inc: the clustering of the extent panner's 1,652
virtual sources (ref src/object_based/polar_extent.cpp:15-39, the same grid for every layout)
into 104 cells of at most 16 points with small bounding caps.

Start: the recursive principal-axis bisection the library builds at run time for any grid
(EARB200_EXTENT_CELLS=bisect).  Then a local search on the sum of the cap radii: neighbouring
cells re-split along a few axes of their union, then single points moved or swapped between
neighbouring cells.  Fewer degrees of radius = fewer (pass, cell) pairs whose sixteen weights
the extent core has to evaluate one by one (tools/exp_cluster.py prints the mixed fraction).
The library checks the table (a permutation of the grid) and computes centres and radii from
the points itself, so the table only decides how good the cells are, never whether the result
is right.  Usage: python tools/gen_extent_cells.py [rounds]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
os.environ["EARB200_EXTENT_CELLS"] = "bisect"
import libear_b200 as eb  # noqa: E402
import exp_cluster as xc  # noqa: E402
import exp_cells as ec  # noqa: E402

CELL = 16


def swap_search(pts, cells, rounds):
    cells = [list(c) for c in cells]
    ctr, rad = xc.caps(pts, cells)
    for rd in range(rounds):
        improved = 0
        for a in np.argsort(-rad):
            pa = pts[cells[a]]
            s_loc = int(np.argmin(pa @ ctr[a]))  # the support point: farthest from the centre
            s = cells[a][s_loc]
            near = [b for b in np.argsort(-(ctr @ ctr[a]))[1:7]]
            best, best_move = 0.0, None
            for b in near:
                base = rad[a] + rad[b]
                cands = []
                if len(cells[b]) < CELL and len(cells[a]) > 1:
                    cands.append(None)  # move
                cands.extend(range(len(cells[b])))  # swaps
                for t_loc in cands:
                    na = [q for q in cells[a] if q != s]
                    nb = list(cells[b])
                    if t_loc is None:
                        nb.append(s)
                    else:
                        na.append(nb[t_loc])
                        nb[t_loc] = s
                    (c1, r1), (c2, r2) = xc.cap(pts[na]), xc.cap(pts[nb])
                    gain = base - (r1 + r2)
                    if gain > 1e-6 and (best_move is None or gain > best):
                        best, best_move = gain, (b, na, nb, c1, r1, c2, r2)
            if best_move:
                b, na, nb, c1, r1, c2, r2 = best_move
                cells[a], cells[b] = na, nb
                ctr[a], rad[a], ctr[b], rad[b] = c1, r1, c2, r2
                improved += 1
        print(f"  swap round {rd}: {improved} moves, mean radius {np.degrees(rad).mean():.3f} "
              f"max {np.degrees(rad).max():.2f}", flush=True)
        if not improved:
            break
    return cells


def cap_ws(p, c0, iters=120):
    """bounding cap from a warm start (Badoiu-Clarkson steps, the best centre seen)"""
    ctr = c0 / np.linalg.norm(c0)
    best_c, best_d = ctr, (p @ ctr).min()
    for it in range(iters):
        j = np.argmin(p @ ctr)
        ctr = ctr + (p[j] - ctr) / (it + 3.0)
        ctr /= np.linalg.norm(ctr)
        d = (p @ ctr).min()
        if d > best_d:
            best_c, best_d = ctr.copy(), d
    return best_c, float(np.arccos(np.clip(best_d, -1, 1)))


def deep_swap(pts, cells, rounds, top=3):
    """swaps of one of the `top` farthest points of a cell with any point of a neighbouring cell
    (and moves into cells with free slots), accepted when the sum of the two radii falls"""
    cells = [list(c) for c in cells]
    ctr, rad = xc.caps(pts, cells)
    for rd in range(rounds):
        improved = 0
        for a in np.argsort(-rad):
            near = [int(b) for b in np.argsort(-(ctr @ ctr[a]))[1:7]]
            far_a = [cells[a][i] for i in np.argsort(pts[cells[a]] @ ctr[a])[:top]]
            best, best_move = 1e-6, None
            for b in near:
                base = rad[a] + rad[b]
                opts = [(s, t) for s in far_a for t in cells[b]]
                if len(cells[b]) < CELL and len(cells[a]) > 1:
                    opts += [(s, None) for s in far_a]
                for s, t in opts:
                    na = [q for q in cells[a] if q != s]
                    nb = [q for q in cells[b] if q != t] + [s]
                    if t is not None:
                        na.append(t)
                    (c1, r1), (c2, r2) = cap_ws(pts[na], ctr[a]), cap_ws(pts[nb], ctr[b])
                    if base - (r1 + r2) > best:
                        # confirm with the full search before accepting
                        (c1, r1), (c2, r2) = xc.cap(pts[na]), xc.cap(pts[nb])
                        if base - (r1 + r2) > best:
                            best, best_move = base - (r1 + r2), (b, na, nb, c1, r1, c2, r2)
            if best_move:
                b, na, nb, c1, r1, c2, r2 = best_move
                cells[a], cells[b] = na, nb
                ctr[a], rad[a], ctr[b], rad[b] = c1, r1, c2, r2
                improved += 1
        print(f"  deep round {rd}: {improved} moves, mean radius {np.degrees(rad).mean():.3f} "
              f"max {np.degrees(rad).max():.2f}", flush=True)
        if not improved:
            break
    return cells


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    grid, order, _, radius = eb.extent_cells(37)
    assert len(grid) == 1652 and order.shape == (104, 16)
    pts = np.asarray(grid, dtype=np.float64)
    cells = [[int(p) for p in row if p >= 0] for row in order]
    rng = np.random.default_rng(2)
    passes = [p for p in ec.pass_params(rng, 400) if max(p[1], p[2]) > 1e-9]
    xc.evaluate(pts, cells, passes, "bisection")
    cells = xc.refine(pts, cells)
    xc.evaluate(pts, cells, passes, "re-split")
    cells = swap_search(pts, cells, rounds)
    xc.evaluate(pts, cells, passes, "moved / swapped")
    if "--deep" in sys.argv:
        cells = deep_swap(pts, cells, rounds)
        xc.evaluate(pts, cells, passes, "deep swaps")
    flat = sorted(p for c in cells for p in c)
    assert flat == list(range(1652)) and len(cells) == 104 and max(len(c) for c in cells) <= CELL
    # cells in the order of the bisection's first point (keeps neighbouring cells close in the list)
    out = os.environ.get("EARB_CELLS_OUT", os.path.join(ROOT, "libear_b200", "csrc", "extent_cells_1652.inc"))
    with open(out, "w") as f:
        f.write("// generated by tools/gen_extent_cells.py -- do not edit.  Cells of the 1,652-point extent grid\n"
                "// (37 rows): 104 x 16 point indices, -1 = empty slot.\n")
        for c in cells:
            row = list(c) + [-1] * (CELL - len(c))
            f.write(" " + ",".join(f"{p:5d}" for p in row) + ",\n")
    print("wrote", out)


if __name__ == "__main__":
    main()

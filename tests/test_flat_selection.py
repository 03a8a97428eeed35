"""The selection rule of the tiny-scene walk (trace_flat, rt_trace.cuh): every sphere's candidate root is computed
against the tmax the walk STARTED with, then the closest is picked in primitive order with the reference's strict
compare -- and that must equal the reference's rule, spheres one after the other with tmax shrinking on every hit
(rt.cu:1094-1121).  Checked exhaustively over small value grids (ties, roots at tmin, misses), no arithmetic involved:
the only fact used is root1 <= root2 (same divisor 2a > 0)."""
import itertools


def sequential(spheres, tmin, tmax):
    best = None
    for i, (valid, r1, r2) in enumerate(spheres):
        if not valid:
            continue
        root = r1
        if tmin >= root or root >= tmax:          # rt.cu:1103-1108
            root = r2
        if not (tmin < root < tmax):
            continue
        tmax, best = root, i
    return best, tmax


def flat(spheres, tmin, tmax0):
    cands = []
    for valid, r1, r2 in spheres:                 # sphere_roots2_fast: against the initial tmax
        root = r2 if (tmin >= r1 or r1 >= tmax0) else r1
        cands.append((valid and tmin < root < tmax0, root))
    best, tbest = None, tmax0
    for i, (hit, root) in enumerate(cands):       # trace_flat: closest in primitive order, strict compare
        if hit and root < tbest:
            tbest, best = root, i
    return best, tbest


def test_flat_selection_equals_shrinking_tmax():
    grid = [0.0, 1.0, 2.0, 3.0, 4.0]              # tmin = 1 sits on the grid: roots at and below tmin are covered
    roots = [(a, b) for a in grid for b in grid if a <= b]
    one = [(False, 2.0, 3.0)] + [(True, a, b) for a, b in roots]
    n = 0
    for tmax0 in (3.0, float("inf")):
        for combo in itertools.product(one, repeat=3):
            assert flat(combo, 1.0, tmax0) == sequential(combo, 1.0, tmax0), (combo, tmax0)
            n += 1
    assert n == 2 * len(one) ** 3

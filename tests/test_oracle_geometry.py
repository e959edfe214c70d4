"""The polygon leaf of the env path (obstacle.py:126-141: shapely ``contains`` / ``exterior.distance``) has no shapely to
be pinned against -- here the restatements (oracle/geometry.py and the map functions of oracle/ship_oracle.c) are checked
against an INDEPENDENT exact evaluator (tests/exact_geometry.py: winding number + exact point-segment distance in
integer / rational arithmetic) on the default map (main_ast.py:200-205): 2*10^5 random grid points, points hugging the
rings, every vertex, edge mid-point, collinear extension, ray-through-vertex and horizontal-edge case."""
import ctypes as C

import numpy as np

import exact_geometry as X
import scenarios as sc
from oracle import cport, geometry

MAP = sc.DEFAULT_MAP


def _c_contains(m, n, e):
    return bool(cport.lib().orc_map_contains(C.byref(m), float(n), float(e)))


def _py_contains(polys, n, e):
    pt = geometry.Point(e, n)
    return any(p.contains(pt) for p in polys)


def _py_distance(polys, n, e):
    pt = geometry.Point(e, n)
    return min(p.exterior.distance(pt) for p in polys)


def test_contains_matches_exact_winding_number_on_random_points():
    rng = np.random.default_rng(25450)
    n = 200_000
    # (x, y) = (east, north) on a 1/1024 grid over the map and a margin around it
    ix = rng.integers(-500 * X.GRID, 10500 * X.GRID, n); iy = rng.integers(-500 * X.GRID, 10500 * X.GRID, n)
    # a quarter of the points hug the rings: a ring point plus a sub-metre offset
    edges = geometry.flatten_edges(MAP)
    for j in range(0, n, 4):
        x1, y1, x2, y2, _ = edges[j % len(edges)]
        t = rng.random()
        ix[j] = int(round((x1 + t * (x2 - x1)) * X.GRID)) + rng.integers(-300, 301)
        iy[j] = int(round((y1 + t * (y2 - y1)) * X.GRID)) + rng.integers(-300, 301)
    inside, bnd = X.locate_int(MAP, ix, iy)
    assert inside.sum() > 0.2 * n and (~inside).sum() > 0.2 * n
    m = cport.make_map(MAP)
    polys = [geometry.Polygon(v) for v in MAP]
    x, y = ix / X.GRID, iy / X.GRID          # exact in double
    got_c = np.array([_c_contains(m, y[i], x[i]) for i in range(n)])
    assert np.array_equal(got_c, inside), f"ship_oracle.c differs from the exact locator at {np.nonzero(got_c != inside)[0][:5]}"
    sub = np.concatenate([np.arange(0, n, 4), np.arange(1, n, 16)])     # the ring-hugging quarter and a sample of the rest
    got_py = np.array([_py_contains(polys, y[i], x[i]) for i in sub])
    assert np.array_equal(got_py, inside[sub]), "oracle/geometry.py differs from the exact locator"


def test_special_points_contains_and_distance():
    m = cport.make_map(MAP)
    polys = [geometry.Polygon(v) for v in MAP]
    pts = X.special_points(MAP)
    assert len(pts) > 400
    n_bnd = 0
    for x, y, tag in pts:
        inside, bnd = X.locate_exact(MAP, x, y)
        n_bnd += bnd
        if tag in ("vertex", "mid", "quarter"):
            assert bnd and not inside, f"{tag} ({x}, {y}) lies on a ring: boundary, hence not contained"
        assert _py_contains(polys, y, x) == inside, f"geometry.py contains at {tag} ({x}, {y})"
        assert _c_contains(m, y, x) == inside, f"ship_oracle.c contains at {tag} ({x}, {y})"
        d = X.sqrt_fraction(X.dist2_exact(MAP, x, y))
        for name, got in (("geometry.py", _py_distance(polys, y, x)), ("ship_oracle.c", cport.lib().orc_map_distance(C.byref(m), float(y), float(x)))):
            assert abs(got - d) <= 1e-12 * max(1.0, d) + (1e-9 if bnd else 0.0), f"{name} distance at {tag} ({x}, {y}): {got} vs {d}"
        if bnd:
            assert d == 0.0
    assert n_bnd >= 3 * 39


def test_distance_matches_exact_on_random_points():
    rng = np.random.default_rng(7)
    m = cport.make_map(MAP)
    polys = [geometry.Polygon(v) for v in MAP]
    worst = 0.0
    for _ in range(3000):
        x, y = float(rng.uniform(-500, 10500)), float(rng.uniform(-500, 10500))
        d = X.sqrt_fraction(X.dist2_exact(MAP, x, y))
        a = _py_distance(polys, y, x)
        b = cport.lib().orc_map_distance(C.byref(m), y, x)
        worst = max(worst, abs(a - d) / max(1.0, d), abs(b - d) / max(1.0, d))
    assert worst <= 1e-12, f"relative distance error {worst:.3e}"


def test_exact_evaluator_on_hand_computed_cases():
    """The checker itself, on cases worked by hand: a 10 x 10 square and a concave 'L'."""
    sq = [[(0, 0), (10, 0), (10, 10), (0, 10)]]
    assert X.locate_exact(sq, 5, 5) == (True, False)
    assert X.locate_exact(sq, 10, 5) == (False, True)            # on an edge
    assert X.locate_exact(sq, 10, 10) == (False, True)           # on a vertex
    assert X.locate_exact(sq, 15, 10) == (False, False)          # level with the top edge, outside
    assert X.locate_exact(sq, -1, 0) == (False, False)           # collinear extension of the bottom edge
    assert X.dist2_exact(sq, 5, 5) == 25 and X.dist2_exact(sq, 13, 14) == 25 and X.dist2_exact(sq, 5, 12) == 4
    ell = [[(0, 0), (4, 0), (4, 2), (2, 2), (2, 4), (0, 4)]]
    assert X.locate_exact(ell, 1, 3) == (True, False) and X.locate_exact(ell, 3, 3) == (False, False)
    assert X.locate_exact(ell, 3, 2) == (False, True) and X.locate_exact(ell, 1, 2) == (True, False)   # ray through the reflex vertex
    assert X.dist2_exact(ell, 3, 3) == 1 and X.dist2_exact(ell, 1, 1) == 1
    ins, bnd = X.locate_int(ell, np.array([1, 3, 3, 1]) * X.GRID, np.array([3, 3, 2, 2]) * X.GRID)
    assert ins.tolist() == [True, False, False, True] and bnd.tolist() == [False, False, True, False]

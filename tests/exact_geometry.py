"""Independent EXACT evaluator for the two polygon queries of the env path (obstacle.py:126-141):
``Polygon.contains(Point)`` (interior only) and ``Polygon.exterior.distance(Point)``.

Written from the mathematics, not from oracle/geometry.py: location is the WINDING NUMBER of the ring about the point
(Sunday's upward / downward crossing rule) with an exact on-segment test first, distance is the exact squared distance to
the closest point of every ring segment.  All arithmetic is exact: integers on a dyadic grid (numpy int64, vectorised)
for location, ``fractions.Fraction`` for distances -- every float is a dyadic rational, so float inputs are exact too.
shapely's semantics (interior only; boundary ring distance) are the documented ones; shapely itself is absent here, so
this pins the restatements to the mathematics rather than to shapely (a21 stays "unpinned vs shapely")."""
from __future__ import annotations

from fractions import Fraction
from math import isqrt

import numpy as np

GRID = 1024   # points for the vectorised integer evaluator are multiples of 1 / GRID


def rings(polygons):
    out = []
    for verts in polygons:
        c = [(x, y) for x, y in verts]
        if len(c) > 1 and c[0] == c[-1]:
            c = c[:-1]
        out.append(c)
    return out


def locate_int(polygons, px, py):
    """px, py: integer arrays (coordinates x GRID).  Returns (inside_any, on_boundary_any) bool arrays, exact."""
    px = np.asarray(px, dtype=np.int64); py = np.asarray(py, dtype=np.int64)
    inside_any = np.zeros(px.shape, bool); bnd_any = np.zeros(px.shape, bool)
    for ring in rings(polygons):
        wn = np.zeros(px.shape, np.int64); bnd = np.zeros(px.shape, bool)
        n = len(ring)
        for i in range(n):
            ax, ay = (int(round(v * GRID)) for v in ring[i])
            bx, by = (int(round(v * GRID)) for v in ring[(i + 1) % n])
            cross = (bx - ax) * (py - ay) - (by - ay) * (px - ax)          # > 0: point left of a -> b
            within = (np.minimum(ax, bx) <= px) & (px <= np.maximum(ax, bx)) & (np.minimum(ay, by) <= py) & (py <= np.maximum(ay, by))
            bnd |= (cross == 0) & within
            up = (ay <= py) & (by > py) & (cross > 0)
            down = (ay > py) & (by <= py) & (cross < 0)
            wn += up.astype(np.int64) - down.astype(np.int64)
        inside_any |= (wn != 0) & ~bnd
        bnd_any |= bnd
    return inside_any, bnd_any


def locate_exact(polygons, x, y):
    """Single point, any float / Fraction coordinates: (inside_any, on_boundary_any)."""
    x, y = Fraction(x), Fraction(y)
    inside_any = bnd_any = False
    for ring in rings(polygons):
        wn, bnd = 0, False
        n = len(ring)
        for i in range(n):
            ax, ay = Fraction(ring[i][0]), Fraction(ring[i][1])
            bx, by = Fraction(ring[(i + 1) % n][0]), Fraction(ring[(i + 1) % n][1])
            cross = (bx - ax) * (y - ay) - (by - ay) * (x - ax)
            if cross == 0 and min(ax, bx) <= x <= max(ax, bx) and min(ay, by) <= y <= max(ay, by):
                bnd = True
            if ay <= y < by and cross > 0:
                wn += 1
            elif by <= y < ay and cross < 0:
                wn -= 1
        inside_any |= (wn != 0 and not bnd)
        bnd_any |= bnd
    return inside_any, bnd_any


def dist2_exact(polygons, x, y) -> Fraction:
    """Exact squared distance from (x, y) to the nearest ring segment of any polygon."""
    x, y = Fraction(x), Fraction(y)
    best = None
    for ring in rings(polygons):
        n = len(ring)
        for i in range(n):
            ax, ay = Fraction(ring[i][0]), Fraction(ring[i][1])
            bx, by = Fraction(ring[(i + 1) % n][0]), Fraction(ring[(i + 1) % n][1])
            dx, dy = bx - ax, by - ay
            l2 = dx * dx + dy * dy
            t = ((x - ax) * dx + (y - ay) * dy) / l2 if l2 else Fraction(0)
            t = min(max(t, Fraction(0)), Fraction(1))
            cx, cy = ax + t * dx, ay + t * dy
            d2 = (x - cx) ** 2 + (y - cy) ** 2
            if best is None or d2 < best:
                best = d2
    return best


def sqrt_fraction(q: Fraction, bits: int = 80) -> float:
    """Correctly rounded-ish sqrt of a non-negative Fraction (error far below a double ulp)."""
    if q == 0:
        return 0.0
    scale = 1 << (2 * bits)
    r = isqrt(q.numerator * scale // q.denominator)
    return float(Fraction(r, 1 << bits))


def special_points(polygons):
    """Vertices, edge mid-points, collinear extensions beyond both ends, points level with every vertex (the ray passes
    through it) on both sides, and points on / beside horizontal segments.  Returns list of (x, y, tag)."""
    pts = []
    for ring in rings(polygons):
        n = len(ring)
        xs = [p[0] for p in ring]
        for i in range(n):
            ax, ay = ring[i]
            bx, by = ring[(i + 1) % n]
            pts.append((ax, ay, "vertex"))
            pts.append(((ax + bx) / 2, (ay + by) / 2, "mid"))
            pts.append((ax + (bx - ax) * 0.25, ay + (by - ay) * 0.25, "quarter"))
            pts.append((bx + (bx - ax) * 0.5, by + (by - ay) * 0.5, "ext+"))
            pts.append((ax - (bx - ax) * 0.5, ay - (by - ay) * 0.5, "ext-"))
            for off in (-3000.0, -0.5, 0.5, 3000.0):
                pts.append((ax + off, ay, "level"))
            if ay == by:
                pts.append((min(ax, bx) - 1.0, ay, "horiz-left")); pts.append((max(ax, bx) + 1.0, ay, "horiz-right"))
            pts.append((min(xs) - 10.0, ay, "far-left")); pts.append((max(xs) + 10.0, ay, "far-right"))
    return pts

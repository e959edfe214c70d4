"""Restatement of the two shapely (GEOS) calls on the env path.  TEST INFRASTRUCTURE.

Reference call sites: ``simulators/ship_in_transit/obstacle.py:107`` (``Polygon(verts)``),
``:128-129`` (``Polygon.contains(Point(e, n))``), ``:140-141``
(``Polygon.exterior.distance(Point(e, n))``).  shapely is an un-vendored, un-pinned
third-party dependency of the reference (absent from all three conda ymls) and is not
installable here, so this file restates the *published* GEOS algorithms:

* ``contains`` -> GEOS ``SimplePointInAreaLocator`` / ``RayCrossingCounter`` (ported from
  JTS): count crossings of the horizontal ray to the right of the point; a point that lies
  on a ring segment is BOUNDARY and therefore *not* contained (interior only).
* ``exterior.distance`` -> GEOS ``Distance::pointToSegment`` minimised over the ring's
  segments: Euclidean distance to the boundary ring, positive for interior points.

PARITY UNPINNED: no shapely exists in this container or in the reference tree to check
against; results are validated on hand-computed cases only (tests/test_oracle_geometry.py).
Vertices are (east, north) = (x, y), as the reference passes them.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple


def _orient(ax, ay, bx, by, px, py) -> int:
    """Sign of the signed area of (a, b, p): +1 = p left of a->b, 0 = collinear."""
    d = (bx - ax) * (py - ay) - (by - ay) * (px - ax)
    return (d > 0) - (d < 0)


def ring_locate(ring: Sequence[Tuple[float, float]], px: float, py: float) -> int:
    """RayCrossingCounter over a closed ring.  Returns 1 interior, 0 boundary, -1 exterior."""
    n = len(ring)
    crossings = 0
    for i in range(n):
        x1, y1 = ring[i]
        x2, y2 = ring[(i + 1) % n]
        if x1 == x2 and y1 == y2:
            continue
        # segment strictly to the left of the point: cannot cross the rightward ray
        if x1 < px and x2 < px:
            continue
        # point coincides with the segment end vertex
        if px == x2 and py == y2:
            return 0
        # horizontal segment at the point's ordinate
        if y1 == py and y2 == py:
            lo, hi = (x1, x2) if x1 <= x2 else (x2, x1)
            if lo <= px <= hi:
                return 0
            continue
        # segment straddles the ray's line (upward edge excludes its top, downward its bottom)
        if (y1 > py and y2 <= py) or (y2 > py and y1 <= py):
            s = _orient(x1, y1, x2, y2, px, py)
            if s == 0:
                return 0
            if y2 < y1:
                s = -s
            if s > 0:
                crossings += 1
    return 1 if (crossings % 2) == 1 else -1


def point_segment_distance(px, py, ax, ay, bx, by) -> float:
    """GEOS Distance::pointToSegment."""
    if ax == bx and ay == by:
        return math.hypot(px - ax, py - ay)
    len2 = (bx - ax) * (bx - ax) + (by - ay) * (by - ay)
    r = ((px - ax) * (bx - ax) + (py - ay) * (by - ay)) / len2
    if r <= 0.0:
        return math.hypot(px - ax, py - ay)
    if r >= 1.0:
        return math.hypot(px - bx, py - by)
    s = ((ay - py) * (bx - ax) - (ax - px) * (by - ay)) / len2
    return abs(s) * math.sqrt(len2)


class Point:
    """Stand-in for ``shapely.geometry.Point`` (obstacle.py:128,140)."""

    def __init__(self, x, y):
        self.x = float(x)
        self.y = float(y)


class _Ring:
    def __init__(self, coords):
        self.coords = coords

    def distance(self, pt: Point) -> float:
        n = len(self.coords)
        best = math.inf
        for i in range(n):
            ax, ay = self.coords[i]
            bx, by = self.coords[(i + 1) % n]
            d = point_segment_distance(pt.x, pt.y, ax, ay, bx, by)
            if d < best:
                best = d
        return best


class Polygon:
    """Stand-in for ``shapely.geometry.Polygon`` (obstacle.py:107): shell only, no holes."""

    def __init__(self, verts):
        coords = [(float(x), float(y)) for x, y in verts]
        if len(coords) > 1 and coords[0] == coords[-1]:
            coords = coords[:-1]
        self._coords = coords
        self.exterior = _Ring(coords)

    def contains(self, pt: Point) -> bool:
        return ring_locate(self._coords, pt.x, pt.y) == 1


def flatten_edges(polygons: List[Sequence[Tuple[float, float]]]):
    """[(x1, y1, x2, y2, poly_id)] for every ring segment of every polygon (closing edge included)."""
    out = []
    for pid, verts in enumerate(polygons):
        coords = [(float(x), float(y)) for x, y in verts]
        if len(coords) > 1 and coords[0] == coords[-1]:
            coords = coords[:-1]
        n = len(coords)
        for i in range(n):
            x1, y1 = coords[i]
            x2, y2 = coords[(i + 1) % n]
            out.append((x1, y1, x2, y2, pid))
    return out

"""
TEST INFRASTRUCTURE (oracle.port): the few planar-geometry operations the reference takes
from shapely / GEOS (absent here), restated.

* ``polygonize`` (voronoi_screen.py:314-319) = bounded faces of the planar graph after
  dangle removal, listed in the order in which GEOS meets their shell rings (directed edges
  in insertion order, forward before reverse, ring = face on the right of the directed
  edge).  That order only matters where two PIL rasters overlap; it cannot be verified
  offline ("parity unpinned").
* ``Polygon.contains`` (voronoi_screen.py:324) / ``prep(poly).disjoint`` (processing_utils.py:
  313-329): crossing-number test with an exact on-segment check in float64.
"""
import math


class Point:
    def __init__(self, *args):
        if len(args) == 1:
            args = tuple(args[0])
        self.x = float(args[0])
        self.y = float(args[1])


class LineString:
    def __init__(self, coords):
        self.coords = [(float(c[0]), float(c[1])) for c in coords]


class _Ring:
    def __init__(self, pts):
        self._pts = pts

    @property
    def xy(self):
        return ([p[0] for p in self._pts], [p[1] for p in self._pts])

    @property
    def coords(self):
        return list(self._pts)


def _orient(ax, ay, bx, by, cx, cy):
    return (bx - ax) * (cy - ay) - (by - ay) * (cx - ax)


class Polygon:
    def __init__(self, verts):
        pts = [(float(v[0]), float(v[1])) for v in verts]
        if pts[0] != pts[-1]:
            pts.append(pts[0])
        self.exterior = _Ring(pts)

    def _locate(self, px, py):
        """+1 interior, 0 boundary, -1 exterior (crossing number + on-segment test)."""
        pts = self.exterior._pts
        inside = False
        for (ax, ay), (bx, by) in zip(pts[:-1], pts[1:]):
            o = _orient(ax, ay, bx, by, px, py)
            if o == 0.0 and min(ax, bx) <= px <= max(ax, bx) and min(ay, by) <= py <= max(ay, by):
                return 0
            if (ay > py) != (by > py):
                # x coordinate of the edge at height py, compared through the orientation sign
                if (o > 0.0) == (by > ay):
                    inside = not inside
        return 1 if inside else -1

    def contains(self, point):
        return self._locate(point.x, point.y) > 0

    def disjoint(self, point):
        return self._locate(point.x, point.y) < 0


class _Prepared:
    def __init__(self, poly):
        self._poly = poly

    def disjoint(self, point):
        return self._poly.disjoint(point)

    def contains(self, point):
        return self._poly.contains(point)


def prep(poly):
    return _Prepared(poly)


def polygonize(lines):
    """Bounded faces of the planar straight-line graph given by 2-point segments."""
    segs = []
    for ln in lines:
        c = ln.coords
        for a, b in zip(c[:-1], c[1:]):
            if a != b:
                segs.append((a, b))
    # adjacency with dangle pruning
    adj = {}
    for a, b in segs:
        adj.setdefault(a, set()).add(b)
        adj.setdefault(b, set()).add(a)
    changed = True
    while changed:
        changed = False
        for v in [v for v, n in adj.items() if len(n) == 1]:
            (u,) = adj[v]
            adj[u].discard(v)
            del adj[v]
            changed = True
        for v in [v for v, n in adj.items() if len(n) == 0]:
            del adj[v]
            changed = True
    # neighbours sorted by angle (counter-clockwise)
    order = {}
    for v, nbrs in adj.items():
        order[v] = sorted(nbrs, key=lambda w: math.atan2(w[1] - v[1], w[0] - v[0]))

    def next_edge(u, v):
        # keep the face on the RIGHT of u->v: at v take the next neighbour counter-clockwise
        # after the reverse direction (v->u), i.e. the sharpest right turn
        nb = order[v]
        i = nb.index(u)
        return v, nb[(i + 1) % len(nb)]

    seen = set()
    polys = []
    for a, b in segs:
        if a not in adj or b not in adj or b not in adj[a]:
            continue
        for start in ((a, b), (b, a)):
            if start in seen:
                continue
            ring = []
            e = start
            while e not in seen:
                seen.add(e)
                ring.append(e[0])
                e = next_edge(*e)
            area2 = 0.0
            for (x0, y0), (x1, y1) in zip(ring, ring[1:] + ring[:1]):
                area2 += x0 * y1 - x1 * y0
            if area2 < 0.0:  # clockwise ring = shell (GEOS: counter-clockwise rings are holes)
                polys.append(Polygon(ring))
    return polys



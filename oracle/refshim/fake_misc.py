"""
TEST INFRASTRUCTURE (oracle): stand-ins for ``shapely``, ``lsmtool`` and the reference's
own ``utils/h5parm.py`` (pytables based; pytables/HDF5 are not installed here).

* shapely is used at voronoi_screen.py:294,314-325 (LineString, polygonize, Polygon,
  contains) and processing_utils.py:313-329 (Polygon, prep().disjoint, Point).
  GEOS semantics restated: ``polygonize`` = bounded faces of the planar graph after dangle
  removal, listed in the order in which GEOS meets their shell rings (directed edges in
  insertion order, forward before reverse, ring = face on the right of the directed
  edge).  That order only matters where two PIL rasters overlap; it cannot be verified
  offline ("parity unpinned").
* lsmtool is used at voronoi_screen.py:241-246 / kl_screen.py:75-80: ``load(f)
  .getPatchPositions()`` -> {patch: [RA, Dec]} with ``.value`` in degrees.
* H5parm: duck-typed in-memory object model with exactly the attributes the path touches
  (stationscreen.py:922-984,1094-1159; kl_screen.py:66-155; voronoi_screen.py:63-102;
  processing_utils.py:557-572; make_aterm_images.py:107-113).
"""
import math

import numpy as np


from oracle.port.geom import (LineString, Point, Polygon, polygonize,  # noqa: F401,E402
                              prep)


# =============================================================================== lsmtool
class _Deg:
    def __init__(self, value):
        self.value = value


def _ra2deg(text):
    h, m, s = text.strip().split(":")
    sign = -1.0 if h.strip().startswith("-") else 1.0
    return sign * (abs(float(h)) + float(m) / 60.0 + float(s) / 3600.0) * 15.0


def _dec2deg(text):
    parts = text.strip().split(".")
    d, m = parts[0], parts[1]
    s = ".".join(parts[2:]) if len(parts) > 2 else "0"
    sign = -1.0 if d.strip().startswith("-") else 1.0
    return sign * (abs(float(d)) + float(m) / 60.0 + float(s) / 3600.0)


class _SkyModel:
    def __init__(self, filename):
        self._patches = {}
        with open(filename, encoding="utf8") as fh:
            for line in fh:
                cols = [c.strip() for c in line.split(",")]
                if len(cols) >= 5 and cols[0] == "" and cols[1] == "" and cols[2]:
                    self._patches[cols[2]] = [_Deg(_ra2deg(cols[3])), _Deg(_dec2deg(cols[4]))]

    def getPatchPositions(self):  # noqa: N802 - lsmtool's name
        return dict(self._patches)


def lsm_load(filename):
    return _SkyModel(filename)

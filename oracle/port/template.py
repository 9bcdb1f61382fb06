"""
TEST INFRASTRUCTURE (oracle.port): restatement of the reference's Voronoi raster template
(voronoi_screen.py:218-351 + processing_utils.py:295-334).

Pipeline: patch RA/Dec -> pixel xy (SIN WCS of the cube header) -> 64 guard sites ->
scipy.spatial.Voronoi -> polygonize finite ridges -> per polygon PIL fill+outline mask with
an exact test of the outline pixels -> later polygons overwrite -> remaining zeros take the
value of the nearest non-zero PIXEL (scipy griddata 'nearest').
"""
import numpy as np
from PIL import Image, ImageDraw
from scipy.interpolate import griddata
from scipy.spatial import Voronoi  # pylint: disable=E0611

from . import geom, wcs


def rasterize(verts, shape):
    """processing_utils.py:295-334 on an all-ones array of ``shape``."""
    data = np.ones(shape)
    poly = geom.Polygon(verts)
    mask = Image.new("L", (shape[0], shape[1]), 0)
    ImageDraw.Draw(mask).polygon(verts, outline=1, fill=1)
    data *= np.array(mask)
    mask = Image.new("L", (shape[0], shape[1]), 0)
    ImageDraw.Draw(mask).polygon(verts, outline=1, fill=0)
    xs, ys = np.where(np.array(mask).transpose())
    for xm, ym in zip(xs, ys):
        if poly.disjoint(geom.Point(xm, ym)):
            data[int(ym), int(xm)] = 0
    return data


def field_corners(ra_deg, dec_deg, rad, dec, width, n, cell):
    """voronoi_screen.py:264-282."""
    bounds = [rad + width / 2.0, dec - width / 2.0, rad - width / 2.0, dec + width / 2.0]
    lo = wcs.voronoi_patch_pixels(max(bounds[0], np.max(ra_deg) + 0.1),
                                  min(bounds[1], np.min(dec_deg) - 0.1), rad, dec, n, cell)
    hi = wcs.voronoi_patch_pixels(min(bounds[2], np.min(ra_deg) - 0.1),
                                  max(bounds[3], np.max(dec_deg) + 0.1), rad, dec, n, cell)
    return (float(lo[0]), float(lo[1])), (float(hi[0]), float(hi[1]))


def guard_points(xy, field_minxy, field_maxxy, nouter=64):
    """voronoi_screen.py:298-309."""
    means = np.ones((nouter, 2)) * np.array(xy).mean(axis=0)
    angles = [np.pi / (nouter / 2.0) * i for i in range(0, nouter)]
    offsets = [[np.cos(a), np.sin(a)] for a in angles]
    radius = 2.0 * np.sqrt((field_maxxy[0] - field_minxy[0]) ** 2
                           + (field_maxxy[1] - field_minxy[1]) ** 2)
    return means + radius * np.array(offsets)


def reference_template(ra_deg, dec_deg, rad, dec, width, cell):
    """Returns (template int32 [ny, nx] of patch index + 1, patch pixel x, patch pixel y)."""
    n = int(np.ceil(width / cell))
    px, py = wcs.voronoi_patch_pixels(ra_deg, dec_deg, rad, dec, n, cell)
    xy = list(zip(px.tolist(), py.tolist()))
    fmin, fmax = field_corners(np.asarray(ra_deg), np.asarray(dec_deg), rad, dec, width, n, cell)
    if len(xy) == 1:
        polygons = [geom.Polygon([fmin, (fmin[0], fmax[1]), fmax, (fmax[0], fmin[1]), fmin])]
    else:
        outer = guard_points(xy, fmin, fmax)
        vor = Voronoi(np.vstack([xy, outer]))
        lines = [geom.LineString(vor.vertices[ln]) for ln in vor.ridge_vertices if -1 not in ln]
        polygons = geom.polygonize(lines)
    for i, p in enumerate(xy):                                    # :322-325
        for poly in polygons:
            if poly.contains(geom.Point(p)):
                poly.index = i
    tmpl = np.zeros((n, n))
    for poly in polygons:                                         # :331-340
        vx, vy = poly.exterior.xy
        raster = rasterize(list(zip(vx, vy)), (n, n)) * (poly.index + 1)
        filled = np.where(raster > 0)
        tmpl[filled] = raster[filled]
    zero = np.where(tmpl == 0)                                    # :341-349
    if len(zero[0]) > 0:
        nz = np.where(tmpl != 0)
        tmpl[zero] = griddata((nz[0], nz[1]), tmpl[nz], (zero[0], zero[1]), method="nearest")
    return tmpl.astype(np.int32), np.asarray(px), np.asarray(py)

"""
Voronoi raster template: which patch (index + 1) every image pixel belongs to.

Two builders, selected by ``VoronoiScreen.template_mode``:

* ``"reference"`` (default, drop-in): the reference's once-per-run HOST geometry pipeline
  (voronoi_screen.py:264-350 + processing_utils.py:295-334): guard ring -> scipy Voronoi ->
  per cell a PIL fill+outline mask whose outline pixels are re-tested exactly -> later
  cells overwrite -> holes take the nearest non-zero pixel.  It is setup, not the data-
  parallel path; scipy / Pillow are the reference's own third-party dependencies.
* ``"nearest"``: the CUDA nearest-patch assignment kernel (``ssf_voronoi_assign``), a pure
  float64 nearest-site map.  It differs from the reference pipeline in O(100) boundary
  pixels per image (PIL integer vertex truncation), see DESIGN.md.
"""
import numpy as np


def field_corners(ra_deg, dec_deg, rad, dec, width, n, cell):
    """Tessellation region in pixels (voronoi_screen.py:264-282)."""
    from . import wcs

    ra_hi = max(rad + width / 2.0, float(np.max(ra_deg)) + 0.1)
    de_lo = min(dec - width / 2.0, float(np.min(dec_deg)) - 0.1)
    ra_lo = min(rad - width / 2.0, float(np.min(ra_deg)) - 0.1)
    de_hi = max(dec + width / 2.0, float(np.max(dec_deg)) + 0.1)
    x0, y0 = wcs.patch_pixels(ra_hi, de_lo, rad, dec, n, cell)
    x1, y1 = wcs.patch_pixels(ra_lo, de_hi, rad, dec, n, cell)
    return (float(x0), float(y0)), (float(x1), float(y1))


def _guard_ring(sites, lo, hi, count=64):
    """64 sites on a circle of radius 2 * field diagonal about the mean patch position
    (voronoi_screen.py:298-309)."""
    centre = np.ones((count, 2)) * sites.mean(axis=0)
    ang = [np.pi / (count / 2.0) * k for k in range(count)]
    unit = np.array([[np.cos(a), np.sin(a)] for a in ang])
    radius = 2.0 * np.sqrt((hi[0] - lo[0]) ** 2 + (hi[1] - lo[1]) ** 2)
    return centre + radius * unit


def _cells_in_polygonize_order(vor, nsites):
    """Bounded cells of the patch sites as vertex rings, listed in the order a polygonizer
    walking the finite ridges (insertion order, each edge forward then backward, face on the
    right) first meets them (voronoi_screen.py:313-319)."""
    order = []
    seen = set()
    verts = vor.vertices
    for (p, q), (a, b) in zip(vor.ridge_points, vor.ridge_vertices):
        if a < 0 or b < 0:
            continue
        va, vb = verts[a], verts[b]
        # which of the two sites lies to the right of a -> b
        cross_p = (vb[0] - va[0]) * (vor.points[p][1] - va[1]) - (vb[1] - va[1]) * (vor.points[p][0] - va[0])
        right, left = (p, q) if cross_p < 0 else (q, p)
        for site in (right, left):
            if site < nsites and site not in seen:
                seen.add(site)
                order.append(int(site))
    rings = []
    for site in order:
        region = vor.regions[vor.point_region[site]]
        if -1 in region or len(region) < 3:
            continue
        pts = verts[region]
        ang = np.arctan2(pts[:, 1] - vor.points[site][1], pts[:, 0] - vor.points[site][0])
        ring = pts[np.argsort(-ang)]                      # clockwise, like a GEOS shell
        rings.append((site, ring))
    return rings


def _strictly_outside_convex(ring, xs, ys):
    """True where integer point (x, y) has no point in common with the convex ring."""
    x0, y0 = ring[:, 0], ring[:, 1]
    x1, y1 = np.roll(x0, -1), np.roll(y0, -1)
    cross = (x1 - x0)[None, :] * (ys[:, None] - y0[None, :]) - (y1 - y0)[None, :] * (xs[:, None] - x0[None, :])
    orient = np.sign(np.sum(x0 * y1 - x1 * y0))           # +1 counter-clockwise
    return np.any(cross * orient < 0.0, axis=1)


def rasterize_cell(ring, shape):
    """processing_utils.py:295-334 on an all-ones image: 1 inside the cell, 0 elsewhere."""
    from PIL import Image, ImageDraw

    verts = [(float(x), float(y)) for x, y in ring]
    verts.append(verts[0])
    full = Image.new("L", (shape[0], shape[1]), 0)
    ImageDraw.Draw(full).polygon(verts, outline=1, fill=1)
    mask = np.array(full).astype(bool)
    edge = Image.new("L", (shape[0], shape[1]), 0)
    ImageDraw.Draw(edge).polygon(verts, outline=1, fill=0)
    ys, xs = np.nonzero(np.array(edge))
    if len(xs):
        out = _strictly_outside_convex(np.asarray(ring, np.float64), xs.astype(np.float64),
                                       ys.astype(np.float64))
        mask[ys[out], xs[out]] = False
    return mask


def reference_template(ra_deg, dec_deg, rad, dec, width, cell):
    """Returns (template int32 [n, n] with patch index + 1, patch pixel x, patch pixel y)."""
    from scipy.interpolate import griddata
    from scipy.spatial import Voronoi  # pylint: disable=no-name-in-module

    from . import wcs

    ra_deg = np.asarray(ra_deg, np.float64)
    dec_deg = np.asarray(dec_deg, np.float64)
    n = wcs.image_size(width, cell)
    px, py = wcs.patch_pixels(ra_deg, dec_deg, rad, dec, n, cell)
    sites = np.column_stack([px, py])
    lo, hi = field_corners(ra_deg, dec_deg, rad, dec, width, n, cell)
    if len(sites) == 1:
        rings = [(0, np.array([lo, (lo[0], hi[1]), hi, (hi[0], lo[1])], np.float64))]
    else:
        vor = Voronoi(np.vstack([sites, _guard_ring(sites, lo, hi)]))
        rings = _cells_in_polygonize_order(vor, len(sites))
    tmpl = np.zeros((n, n), np.int32)
    for site, ring in rings:
        tmpl[rasterize_cell(ring, (n, n))] = site + 1
    holes = np.nonzero(tmpl == 0)
    if len(holes[0]):
        known = np.nonzero(tmpl)
        if not len(known[0]):
            raise ValueError("Voronoi template is empty: no patch cell intersects the image")
        tmpl[holes] = griddata((known[0], known[1]), tmpl[known], (holes[0], holes[1]),
                               method="nearest")
    return tmpl, np.asarray(px), np.asarray(py)

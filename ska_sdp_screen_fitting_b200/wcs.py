"""
Closed-form celestial projections used by the hot path (no astropy / wcslib on the image).

The reference builds ``astropy.wcs.WCS`` objects and calls ``wcs_world2pix`` /
``wcs_pix2world``; for its two projections those are the zenithal gnomonic (TAN) and
orthographic (SIN) formulas of FITS WCS paper II, evaluated here directly in float64:

* screen coordinates of pierce points and image pixels (TAN about a refined mid point,
  0.0005 deg "pixels", CRPIX 1000)        stationscreen.py:138-300
* diagonal-pixel RA/Dec of the KL image   kl_screen.py:238-259
* patch pixel positions in the cube WCS   voronoi_screen.py:251-261 (SIN,
  header of processing_utils.py:206-218)
"""
import numpy as np

_RAD = np.pi / 180.0
SCREEN_CRPIX = 1000.0        # stationscreen.py:294
SCREEN_CDELT = 0.0005        # stationscreen.py:295 (x axis negative)


def _std_coords(ra_deg, dec_deg, ra0_deg, dec0_deg, gnomonic):
    """Standard coordinates (xi, eta) in DEGREES of (ra, dec) about (ra0, dec0)."""
    ra = np.asarray(ra_deg, np.float64) * _RAD
    de = np.asarray(dec_deg, np.float64) * _RAD
    ra0, de0 = float(ra0_deg) * _RAD, float(dec0_deg) * _RAD
    dra = ra - ra0
    sd, cd, sd0, cd0 = np.sin(de), np.cos(de), np.sin(de0), np.cos(de0)
    xi = cd * np.sin(dra)
    eta = sd * cd0 - cd * sd0 * np.cos(dra)
    if gnomonic:
        cosc = sd * sd0 + cd * cd0 * np.cos(dra)
        xi, eta = xi / cosc, eta / cosc
    return xi / _RAD, eta / _RAD


def sky_to_pixel(ra_deg, dec_deg, crval, crpix, cdelt, projection, origin=0):
    """``wcs_world2pix(..., origin)`` of a two-axis RA/Dec WCS, projection TAN or SIN."""
    if projection not in ("TAN", "SIN"):
        raise ValueError(f"unsupported projection {projection}")
    xi, eta = _std_coords(ra_deg, dec_deg, crval[0], crval[1], projection == "TAN")
    shift = 1 - origin
    return xi / cdelt[0] + crpix[0] - shift, eta / cdelt[1] + crpix[1] - shift


def pixel_to_sky_tan(x, y, crval, crpix, cdelt, origin=0):
    """``wcs_pix2world(..., origin)`` of a RA---TAN / DEC--TAN WCS; RA wrapped to [0, 360)."""
    shift = 1 - origin
    xi = (np.asarray(x, np.float64) + shift - crpix[0]) * cdelt[0] * _RAD
    eta = (np.asarray(y, np.float64) + shift - crpix[1]) * cdelt[1] * _RAD
    ra0, de0 = float(crval[0]) * _RAD, float(crval[1]) * _RAD
    den = np.cos(de0) - eta * np.sin(de0)
    ra = ra0 + np.arctan2(xi, den)
    de = np.arctan2(np.sin(de0) + eta * np.cos(de0), np.hypot(xi, den))
    return np.mod(ra / _RAD, 360.0), de / _RAD


def screen_xy(ra_deg, dec_deg, ref_ra, ref_dec):
    """_radec2xy with _make_wcs (stationscreen.py:189-231, 275-300)."""
    return sky_to_pixel(ra_deg, dec_deg, (ref_ra, ref_dec), (SCREEN_CRPIX, SCREEN_CRPIX),
                        (-SCREEN_CDELT, SCREEN_CDELT), "TAN", origin=0)


def refined_screen_xy(ra_deg, dec_deg, mid_ra=None, mid_dec=None):
    """_getxy (stationscreen.py:138-186).  Without a given mid point: project about source 0,
    take the first source beyond the x (resp. y) midpoint in sorted order as the new reference
    RA (resp. Dec), and project again.  Returns (x, y, mid_ra, mid_dec)."""
    ra = np.atleast_1d(np.asarray(ra_deg, np.float64))
    de = np.atleast_1d(np.asarray(dec_deg, np.float64))
    if mid_ra is None or mid_dec is None:
        mid_ra, mid_dec = ra[0], de[0]
        if ra.size > 1:
            x0, y0 = screen_xy(ra, de, ra[0], de[0])
            half_x = x0.min() + (x0.max() - x0.min()) / 2.0
            half_y = y0.min() + (y0.max() - y0.min()) / 2.0
            order_x, order_y = np.argsort(x0), np.argsort(y0)
            beyond_x = np.nonzero(x0[order_x] > half_x)[0]
            beyond_y = np.nonzero(y0[order_y] > half_y)[0]
            if beyond_x.size and beyond_y.size:
                mid_ra = ra[order_x[beyond_x[0]]]
                mid_dec = de[order_y[beyond_y[0]]]
    x, y = screen_xy(ra, de, mid_ra, mid_dec)
    return x, y, float(mid_ra), float(mid_dec)


def image_size(width_deg, cellsize_deg):
    """screen.py:180-181 / kl_screen.py:238-239."""
    return int(np.ceil(width_deg / cellsize_deg))


def kl_grid(rad, dec, width_deg, cellsize_deg, mid_ra, mid_dec):
    """Screen coordinates of the KL image grid (kl_screen.py:238-259): RA/Dec of the image
    DIAGONAL pixels (i, i) through a TAN WCS (crpix n/2, cdelt -/+cell), projected about
    (mid_ra, mid_dec).  The image grid is the outer product X[i] x Y[j]."""
    n = image_size(width_deg, cellsize_deg)
    idx = np.arange(n, dtype=np.float64)
    ra, de = pixel_to_sky_tan(idx, idx, (rad, dec), (n / 2.0, n / 2.0),
                              (-cellsize_deg, cellsize_deg), origin=0)
    x, y, _, _ = refined_screen_xy(ra, de, mid_ra, mid_dec)
    return x, y


def patch_pixels(ra_deg, dec_deg, rad, dec, n, cellsize_deg):
    """Patch positions in image pixels through the cube's SIN WCS (voronoi_screen.py:251-261)."""
    return sky_to_pixel(ra_deg, dec_deg, (rad, dec), (n / 2.0, n / 2.0),
                        (-cellsize_deg, cellsize_deg), "SIN", origin=0)

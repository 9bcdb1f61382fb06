"""
TEST INFRASTRUCTURE (oracle.port): closed-form celestial projections.

The reference delegates these to astropy.wcs / wcslib (absent here); the formulas are the
zenithal TAN / SIN projections of FITS WCS paper II.
Parity with wcslib is "unpinned" at the bit level; the same formulas were checked against
the reference's own e2e assertion (tests/test_fit_screens.py:190-215) during the survey.
"""
import numpy as np

D2R = np.pi / 180.0
R2D = 180.0 / np.pi


def world2pix(ra_deg, dec_deg, crval, crpix, cdelt, proj, origin=0):
    """astropy ``wcs_world2pix(..., origin)`` for a 2-axis RA/DEC ``proj`` in {TAN, SIN}."""
    a = np.asarray(ra_deg, dtype=np.float64) * D2R
    d = np.asarray(dec_deg, dtype=np.float64) * D2R
    a0 = crval[0] * D2R
    d0 = crval[1] * D2R
    da = a - a0
    xi = np.cos(d) * np.sin(da)
    eta = np.sin(d) * np.cos(d0) - np.cos(d) * np.sin(d0) * np.cos(da)
    if proj == "TAN":
        den = np.sin(d) * np.sin(d0) + np.cos(d) * np.cos(d0) * np.cos(da)
        xi = xi / den
        eta = eta / den
    elif proj != "SIN":
        raise NotImplementedError(proj)
    x = R2D * xi / cdelt[0] + crpix[0] - (1 - origin)
    y = R2D * eta / cdelt[1] + crpix[1] - (1 - origin)
    return x, y


def pix2world_tan(x, y, crval, crpix, cdelt, origin=0):
    """astropy ``wcs_pix2world(..., origin)`` for RA---TAN / DEC--TAN."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64)
    a0 = crval[0] * D2R
    d0 = crval[1] * D2R
    xi = cdelt[0] * (x + (1 - origin) - crpix[0]) * D2R
    eta = cdelt[1] * (y + (1 - origin) - crpix[1]) * D2R
    den = np.cos(d0) - eta * np.sin(d0)
    a = a0 + np.arctan2(xi, den)
    d = np.arctan2(np.sin(d0) + eta * np.cos(d0), np.hypot(xi, den))
    return (a * R2D) % 360.0, d * R2D


# ------------------------------------------------------------------ screen coordinates
_SCREEN_CRPIX = (1000.0, 1000.0)          # stationscreen.py:294
_SCREEN_CDELT = (-0.0005, 0.0005)         # stationscreen.py:295


def radec2xy(ra_deg, dec_deg, ref_ra=None, ref_dec=None):
    """stationscreen.py:189-231 (_radec2xy + _make_wcs:275-300)."""
    ra_deg = np.asarray(ra_deg, dtype=np.float64)
    dec_deg = np.asarray(dec_deg, dtype=np.float64)
    if ref_ra is None:
        ref_ra = ra_deg[0]
    if ref_dec is None:
        ref_dec = dec_deg[0]
    return world2pix(ra_deg, dec_deg, (ref_ra, ref_dec), _SCREEN_CRPIX, _SCREEN_CDELT, "TAN", 0)


def getxy(ra_deg, dec_deg, mid_ra=None, mid_dec=None):
    """stationscreen.py:138-186 (_getxy): midpoint-refined reference, then projection."""
    ra_deg = np.asarray(ra_deg, dtype=np.float64)
    dec_deg = np.asarray(dec_deg, dtype=np.float64)
    if mid_ra is None or mid_dec is None:
        x, y = radec2xy(ra_deg, dec_deg)
        if len(x) > 1:
            xmid = x.min() + (x.max() - x.min()) / 2.0
            ymid = y.min() + (y.max() - y.min()) / 2.0
            xind = np.argsort(x)
            yind = np.argsort(y)
            wx = np.where(x[xind] > xmid)[0]
            wy = np.where(y[yind] > ymid)[0]
            if len(wx) and len(wy):
                mid_ra = ra_deg[xind[wx[0]]]
                mid_dec = dec_deg[yind[wy[0]]]
            else:
                mid_ra, mid_dec = ra_deg[0], dec_deg[0]
        else:
            mid_ra, mid_dec = ra_deg[0], dec_deg[0]
    x, y = radec2xy(ra_deg, dec_deg, mid_ra, mid_dec)
    return np.array([x, y]), float(mid_ra), float(mid_dec)


def kl_grid_coords(rad, dec, width_deg, cellsize_deg, mid_ra, mid_dec):
    """kl_screen.py:238-259: screen coordinates of the image DIAGONAL pixels (i, i)."""
    n = int(np.ceil(width_deg / cellsize_deg))
    idx = np.arange(n, dtype=np.float64)
    ra, de = pix2world_tan(idx, idx, (rad, dec), (n / 2.0, n / 2.0),
                           (-cellsize_deg, cellsize_deg), 0)
    xy, _, _ = getxy(ra, de, mid_ra, mid_dec)
    return xy[0], xy[1]


def voronoi_patch_pixels(ra_deg, dec_deg, rad, dec, n, cellsize_deg):
    """voronoi_screen.py:251-261 through the cube header's SIN WCS (processing_utils.py:206-218)."""
    return world2pix(ra_deg, dec_deg, (rad, dec), (n / 2.0, n / 2.0),
                     (-cellsize_deg, cellsize_deg), "SIN", 0)

"""
Public API -- drop-in for the reference's ``make_aterm_images.py:15-153``: same name, same
keyword arguments, same files written (``{outroot}_{g}.fits``, ``{outroot}.txt`` and, for
tessellated screens, ``{outroot}_template.fits``).
"""
import os

from .h5parm_io import count_directions
from .kl_screen import KLScreen
from .voronoi_screen import VoronoiScreen


def _float_list(value):
    """Lists may arrive as "[a;b;c;d]" strings from the CLI (make_aterm_images.py:80-87)."""
    if isinstance(value, str):
        return [float(item.strip()) for item in value.strip("[]").split(";")]
    return value


def make_aterm_image(h5parmfile, soltabname="phase000", screen_type="tessellated", outroot="",
                     bounds_deg=None, bounds_mid_deg=None, skymodel=None, solsetname="sol000",
                     padding_fraction=1.4, cellsize_deg=0.2, smooth_deg=0, ncpu=0):
    """Make a-term FITS images from the solutions of ``h5parmfile``.

    ``bounds_deg`` = [maxRA, minDec, minRA, maxDec], ``bounds_mid_deg`` = [RA, Dec] of the
    image centre; ``screen_type`` is "tessellated" (Voronoi) or "kl".  ``ncpu`` is accepted for
    compatibility and ignored: the work runs on the GPU.  Returns None, like the reference
    (make_aterm_images.py:146-153)."""
    if "gain" in soltabname:                                   # :71-78
        soltab_amp = soltabname.replace("gain", "amplitude")
        soltab_ph = soltabname.replace("gain", "phase")
    else:
        soltab_amp = None
        soltab_ph = soltabname
    bounds_deg = _float_list(bounds_deg)
    bounds_mid_deg = _float_list(bounds_mid_deg)
    if bounds_deg is None or bounds_mid_deg is None:
        raise TypeError("bounds_deg and bounds_mid_deg are required")
    if padding_fraction is not None:                           # :88-97
        padding_fraction = float(padding_fraction)
        pad_ra = (bounds_deg[2] - bounds_deg[0]) * (padding_fraction - 1.0)
        pad_dec = (bounds_deg[3] - bounds_deg[1]) * (padding_fraction - 1.0)
        bounds_deg[0] -= pad_ra
        bounds_deg[1] -= pad_dec
        bounds_deg[2] += pad_ra
        bounds_deg[3] += pad_dec
    cellsize_deg = float(cellsize_deg)
    smooth_pix = float(smooth_deg) / cellsize_deg              # :98-100
    if screen_type == "kl":
        smooth_pix = 0.0                                       # :101-103
    if count_directions(h5parmfile, solsetname, soltab_ph) == 1:
        screen_type = "tessellated"                            # :107-113
    width_deg = bounds_deg[3] - bounds_deg[1]                  # :116-118
    rootname = os.path.basename(outroot)
    classes = {"kl": KLScreen, "tessellated": VoronoiScreen}
    if screen_type not in classes:
        raise ValueError(f"unknown screen_type {screen_type!r} (expected 'kl' or 'tessellated')")
    screen = classes[screen_type](rootname, h5parmfile, skymodel, bounds_mid_deg[0],
                                  bounds_mid_deg[1], width_deg, width_deg,
                                  solset_name=solsetname, phase_soltab_name=soltab_ph,
                                  amplitude_soltab_name=soltab_amp)
    screen.process(ncpu=ncpu)
    screen.write(os.path.dirname(outroot), cellsize_deg, smooth_pix=smooth_pix, ncpu=ncpu)

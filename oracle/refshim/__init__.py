"""
TEST INFRASTRUCTURE (oracle): run the reference's OWN modules, unmodified, under a
compatibility shim -- from ``/root/reference/src`` in the build container, or from the copy
``oracle/build_ref.py`` leaves in ``oracle/_ref`` (git-ignored, ships to the GPU box) elsewhere.

Used by ``oracle/make_golden*.py`` to generate ``tests/golden/*.npz``, by the ``not gpu`` tests
that pin ``oracle.port`` against the reference (skipped when no reference tree is available),
and by ``bench.py``'s CPU arm (``oracle/refrun.py``).

What is shimmed and why:
  astropy.wcs / astropy.io.fits / astropy.coordinates  -> fake_astropy  (not installed)
  shapely.geometry / shapely.ops / shapely.prepared     -> fake_misc     (not installed)
  lsmtool                                               -> fake_misc     (not installed)
  ska_sdp_screen_fitting.utils.h5parm                   -> fake_h5parm   (needs pytables)
  numpy.float                                           -> float         (removed in numpy 1.24)
  scipy.linalg.pinv(rcond=)                             -> pinv(atol=rcond, rtol=0), the
        mapping scipy 1.7.3 (the reference's pin, poetry.lock:450-451) applied itself.
"""
import importlib
import os
import sys
import types

_MOUNT = "/root/reference/src"
_COPY = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_ref")
REFERENCE_SRC = _MOUNT if os.path.isdir(os.path.join(_MOUNT, "ska_sdp_screen_fitting")) else _COPY
_LOADED = {}


def available():
    return os.path.isfile(os.path.join(REFERENCE_SRC, "ska_sdp_screen_fitting", "stationscreen.py"))


def _module(name, **attrs):
    mod = types.ModuleType(name)
    mod.__dict__.update(attrs)
    sys.modules[name] = mod
    return mod


def install():
    """Registers the stub modules and imports the reference.  Idempotent."""
    if _LOADED:
        return _LOADED
    if not available():
        raise RuntimeError("reference tree not present; refshim is container-only")
    import numpy as np
    import scipy.linalg

    from . import fake_astropy, fake_h5parm, fake_misc

    sys.dont_write_bytecode = True
    if not hasattr(np, "float"):
        np.float = float  # processing_utils.py:569

    wcs_mod = _module("astropy.wcs", WCS=fake_astropy.WCS)
    fits_mod = _module(
        "astropy.io.fits",
        PrimaryHDU=fake_astropy.PrimaryHDU,
        HDUList=fake_astropy.HDUList,
        Header=fake_astropy.Header,
        open=fake_astropy.open,
    )
    io_mod = _module("astropy.io", fits=fits_mod)
    coord_mod = _module("astropy.coordinates", Angle=fake_astropy.Angle)
    _module("astropy", wcs=wcs_mod, io=io_mod, coordinates=coord_mod)

    geom = _module("shapely.geometry", Point=fake_misc.Point, Polygon=fake_misc.Polygon,
                   LineString=fake_misc.LineString)
    ops = _module("shapely.ops", polygonize=fake_misc.polygonize)
    prepared = _module("shapely.prepared", prep=fake_misc.prep)
    _module("shapely", geometry=geom, ops=ops, prepared=prepared)
    _module("lsmtool", load=fake_misc.lsm_load)

    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    _module("ska_sdp_screen_fitting.utils.h5parm", H5parm=fake_h5parm.H5parm,
            Solset=fake_h5parm.Solset, Soltab=fake_h5parm.Soltab)

    def pinv_compat(a, rcond=None, **kw):
        if rcond is not None:
            return scipy.linalg.pinv(a, atol=rcond, rtol=0.0, **kw)
        return scipy.linalg.pinv(a, **kw)

    names = ["stationscreen", "screen", "kl_screen", "voronoi_screen", "make_aterm_images",
             "utils.processing_utils"]
    for n in names:
        _LOADED[n.split(".")[-1]] = importlib.import_module(f"ska_sdp_screen_fitting.{n}")
    _LOADED["stationscreen"].pinv = pinv_compat
    _LOADED["fake_h5parm"] = fake_h5parm
    return _LOADED

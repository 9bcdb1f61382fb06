"""
TEST INFRASTRUCTURE (oracle): drive the reference's OWN public API (``make_aterm_image``,
make_aterm_images.py:15-153 -> ``Screen.process`` / ``Screen.write``) on synthetic solution sets.

The reference reads an h5parm file and a sky-model text file.  There is no HDF5 library here, so
the synthetic soltab is registered with the in-memory h5parm of ``refshim/fake_h5parm.py``
(``add_virtual`` / ``add_soltab``) under a file name, and the patch positions are written as a
real sky-model file in the format of ``resources/skymodel.txt``.  Everything between those two
inputs and the FITS files written is the reference's unmodified code (``oracle.refshim``).

Used by ``bench.py`` (``--impl reference``, ``cpu_baseline``: ``kind: "reference"``) and by the
golden-vector generators.
"""
import os

import numpy as np


def _hms(ra_deg):
    h = (ra_deg % 360.0) / 15.0
    hh = int(h)
    mm = int((h - hh) * 60.0)
    ss = (h - hh - mm / 60.0) * 3600.0
    return f"{hh}:{mm:02d}:{ss:013.10f}"


def _dms(dec_deg):
    sign = "-" if dec_deg < 0 else ""
    d = abs(dec_deg)
    dd = int(d)
    mm = int((d - dd) * 60.0)
    ss = (d - dd - mm / 60.0) * 3600.0
    return f"{sign}{dd}.{mm:02d}.{ss:013.10f}"


def write_skymodel(path, names, radec_deg):
    """Patch lines of an LSMTool sky model (resources/skymodel.txt:1-13: RA h:m:s, Dec d.m.s)."""
    with open(path, "w", encoding="utf8") as fh:
        fh.write("FORMAT = Name, Type, Patch, Ra, Dec, I, SpectralIndex='[]', LogarithmicSI, "
                 "ReferenceFrequency='150616963.704427', MajorAxis, MinorAxis, Orientation\n\n")
        for name, (ra, dec) in zip(names, radec_deg):
            fh.write(f" , , {name.strip('[]')}, {_hms(float(ra))}, {_dms(float(dec))}\n")


def register(sol, h5name, soltab_name="phase000", solset="sol000", amp=None, amp_name="amplitude000",
             soltype="phase"):
    """Makes ``sol`` (a product ``Solutions`` record, or anything with the same fields) readable
    by the reference as soltab ``soltab_name`` of the in-memory h5parm ``h5name``."""
    from oracle.refshim import fake_h5parm

    fake_h5parm.add_virtual(h5name, solset, {n: np.asarray(p, np.float32) for n, p in sol.station_dict.items()},
                            {n: np.asarray(p, np.float32) for n, p in sol.source_dict.items()})
    fake_h5parm.add_soltab(h5name, solset, soltab_name, soltype, ["time", "freq", "ant", "dir"],
                           [np.asarray(sol.time, float), np.asarray(sol.freq, float), np.array(list(sol.ant)),
                            np.array(list(sol.dir))], np.asarray(sol.val), np.asarray(sol.weight))
    if amp is not None:
        fake_h5parm.add_soltab(h5name, solset, amp_name, "amplitude", ["time", "freq", "ant", "dir", "pol"],
                               [np.asarray(amp.time, float), np.asarray(amp.freq, float), np.array(list(amp.ant)),
                                np.array(list(amp.dir)), np.array(["XX", "YY"])], np.asarray(amp.val),
                               np.asarray(amp.weight))


def run_case(case, screen_type, out_dir, smooth_deg=0.0, ncpu=0, name=None, soltabname="phase000", amp=None):
    """The reference's ``make_aterm_image`` on a synthetic ``Case`` (tests/synthetic_cases.py).
    Returns (list of FITS files, dict of the in-memory soltabs the fit created)."""
    from oracle import refshim
    from oracle.refshim import fake_h5parm

    mods = refshim.install()
    name = name or f"ref_{screen_type}"
    h5name = os.path.join(out_dir, f"{name}.h5")
    sky = os.path.join(out_dir, f"{name}_skymodel.txt")
    sol = case.solutions
    register(sol, h5name, amp=amp)
    write_skymodel(sky, sol.dir, case.patch_radec)
    half = case.width_deg / 2.0
    bounds = [case.rad + half, case.dec - half, case.rad - half, case.dec + half]
    mods["make_aterm_images"].make_aterm_image(
        h5name, soltabname=soltabname, screen_type=screen_type, outroot=os.path.join(out_dir, name),
        bounds_deg=bounds, bounds_mid_deg=[case.rad, case.dec], skymodel=sky, solsetname="sol000",
        padding_fraction=None, cellsize_deg=case.cellsize_deg, smooth_deg=smooth_deg, ncpu=ncpu)
    with open(os.path.join(out_dir, f"{name}.txt"), encoding="utf8") as fh:
        files = [ln.strip() for ln in fh if ln.strip()]
    created = fake_h5parm._CREATED.get(h5name, {}).get("sol000", {})
    return files, {k: v._soltab for k, v in created.items()}

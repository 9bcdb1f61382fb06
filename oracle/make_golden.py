"""
TEST INFRASTRUCTURE (oracle): generate ``tests/golden/*.npz`` by running the REFERENCE's own
code (unmodified, under ``oracle.refshim``) on the repo inputs and on small seeded
synthetic cases.  Run in the build container only:

    python -m oracle.make_golden [--only voronoi|kl|fit|smooth]

The fixtures are committed; this script is the provenance record for them.
Cube fixtures keep a sub-sample of (freq, station) pairs to stay small:
freqs (0, 5, 11) x stations (0, 1, 30, 61), all 20 times, all 4 planes, all pixels.
"""
import argparse
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")
SUB_F = (0, 5, 11)
SUB_A = (0, 1, 30, 61)

TEST_ARGS = dict(  # tests/test_fit_screens.py:50-63 and :139-152
    soltabname="phase000",
    bounds_deg=[124.565, 66.165, 127.895, 62.835],
    bounds_mid_deg=[126.23, 64.50],
    skymodel="skymodel.txt",
    solsetname="sol000",
    padding_fraction=0,
    cellsize_deg=0.2,
    ncpu=0,
)


def _workdir():
    tmp = tempfile.mkdtemp(prefix="ssf_golden_")
    for f in ("solutions.h5", "skymodel.txt"):
        shutil.copy(os.path.join(GOLDEN, f), tmp)
    return tmp


def _sub(cube):
    return cube[:, SUB_F][:, :, SUB_A].copy()


def golden_voronoi(mods):
    from oracle.refshim import fake_astropy

    tmp = _workdir()
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        out = {}
        for tag, smooth in (("smooth", 0.1), ("nosmooth", 0.0)):
            args = dict(TEST_ARGS, bounds_deg=list(TEST_ARGS["bounds_deg"]))
            mods["make_aterm_images"].make_aterm_image(
                "solutions.h5", screen_type="tessellated", outroot="tessellated",
                smooth_deg=smooth, **args)
            hdu = fake_astropy.open("tessellated_0.fits")
            out[f"cube_{tag}"] = _sub(hdu[0].data)
            out["header_keys"] = np.array([k for k, _ in hdu[0].header.items()])
            out["header_vals"] = np.array([str(v) for _, v in hdu[0].header.items()])
        # the template itself (values = patch index + 1), via a fresh object
        scr = mods["voronoi_screen"].VoronoiScreen(
            "tessellated", "solutions.h5", "skymodel.txt", 126.23, 64.50, 3.33, 3.33)
        scr.process()
        scr.make_rasertize_template(0.2, ".")
        out["template"] = scr.data_rasertize_template.astype(np.int32)
        out["poly_index"] = np.array([p.index for p in scr.polygons], dtype=np.int32)
        with open("tessellated.txt", encoding="utf8") as fh:
            out["filelist"] = np.array(fh.read().split())
        np.savez_compressed(os.path.join(GOLDEN, "voronoi_c1.npz"), sub_f=SUB_F, sub_a=SUB_A, **out)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)


def golden_kl(mods):
    from oracle.refshim import fake_astropy

    tmp = _workdir()
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        args = dict(TEST_ARGS, bounds_deg=list(TEST_ARGS["bounds_deg"]))
        mods["make_aterm_images"].make_aterm_image(
            "solutions.h5", screen_type="kl", outroot="kl", smooth_deg=0.1, **args)
        hdu = fake_astropy.open("kl_0.fits")
        store = mods["fake_h5parm"]._CREATED["solutions.h5"]["sol000"]
        scr = store["phase_screen000"]._soltab
        res = store["phase_screen000resid"]._soltab
        np.savez_compressed(
            os.path.join(GOLDEN, "kl_c2.npz"),
            sub_f=SUB_F, sub_a=SUB_A,
            cube=_sub(hdu[0].data),
            screen=scr._val, screen_weights=scr._weight,
            resid=res._val, order=res._weight[..., 0],
            piercepoints=np.array(store["phase_screen000"].piercepoint),
            midra=scr.obj._v_attrs["midra"], middec=scr.obj._v_attrs["middec"],
            beta=scr.obj._v_attrs["beta"], r_0=scr.obj._v_attrs["r_0"],
        )
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default=None)
    ns = ap.parse_args()
    sys.path.insert(0, ROOT)
    from oracle import refshim

    mods = refshim.install()
    todo = {"voronoi": golden_voronoi, "kl": golden_kl}
    try:
        from oracle import make_golden_synth

        todo.update(make_golden_synth.JOBS)
    except ImportError:
        pass
    for name, fn in todo.items():
        if ns.only in (None, name):
            print("golden:", name, flush=True)
            fn(mods)


if __name__ == "__main__":
    main()

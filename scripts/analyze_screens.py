"""Screens against the input solutions at the patch positions -- the numeric stand-in for the
reference's plotting script (/root/reference/scripts/analyze_screens.py:97-223), which overlays
the calibration solutions on the KL and Voronoi screens with matplotlib (not in this image).

Same inputs, same steps: the patch positions of the sky model become pixel coordinates of the
cube's own WCS (STEP 2), the solution phases are referenced to antenna 0 (STEP 4,
get_phase_corrected), the expected plane value at a patch is cos / sin of that phase times the
amplitude (STEP 5) -- and instead of colouring a scatter plot with it, the value of each screen at
the patch's pixel is compared with it, over every (time, freq, antenna) or one selected cell.

    python scripts/analyze_screens.py solutions.h5 skymodel.txt kl_0.fits [tessellated_0.fits]
           [--soltab phase000] [--plane 1] [--time T --freq F --antenna A] [--dump out.npz]

Prints per screen: patches inside the image, max / rms of (screen - solution) over all cells, the
worst cell, and -- with a cell selected -- the per-patch table the reference's figure shows.
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("solutions")
    ap.add_argument("skymodel")
    ap.add_argument("cubes", nargs="+", help="a-term cubes written by make_aterm_image (kl_0.fits, tessellated_0.fits ...)")
    ap.add_argument("--solset", default="sol000")
    ap.add_argument("--soltab", default="phase000")
    ap.add_argument("--plane", type=int, default=1, help="MATRIX plane: 0 Re XX, 1 Im XX, 2 Re YY, 3 Im YY")
    ap.add_argument("--time", type=int, default=None)
    ap.add_argument("--freq", type=int, default=None)
    ap.add_argument("--antenna", type=int, default=None)
    ap.add_argument("--ref-antenna", type=int, default=0)
    ap.add_argument("--dump", default=None, help="write the sampled values to this .npz")
    args = ap.parse_args()
    from ska_sdp_screen_fitting_b200 import fits, skymodel, wcs
    from ska_sdp_screen_fitting_b200.h5parm_io import read_solutions

    sol = read_solutions(args.solutions, args.solset, args.soltab)
    phase = np.asarray(sol.val, np.float64)                           # [t, f, a, d]
    positions = skymodel.patch_positions(args.skymodel)
    radec = np.array([positions[str(name).strip("[]")] for name in sol.dir])
    # STEP 4: reference the phases (analyze_screens.py:69-94)
    corrected = phase - phase[:, :, args.ref_antenna:args.ref_antenna + 1, :]
    func = (np.cos, np.sin)[args.plane % 2]
    expected = func(corrected)                                        # amplitudes are ones (:199)
    dump = {"expected": expected, "radec": radec}
    status = 0
    for path in args.cubes:
        header, data = fits.read(path)
        if data.ndim != 6:
            print(f"{path}: {data.ndim}-axis cube (a dTEC cube has no MATRIX axis): skipped")
            continue
        n_t, n_f, n_a, _, n_y, n_x = data.shape
        # STEP 2: patch pixels through the header's own WCS (RA---SIN / DEC--SIN, :152-156)
        px, py = wcs.sky_to_pixel(radec[:, 0], radec[:, 1], (header["CRVAL1"], header["CRVAL2"]),
                                  (header["CRPIX1"], header["CRPIX2"]), (header["CDELT1"], header["CDELT2"]),
                                  "SIN", origin=0)
        ix, iy = np.rint(px).astype(int), np.rint(py).astype(int)
        inside = (ix >= 0) & (ix < n_x) & (iy >= 0) & (iy < n_y)      # STEP 3
        if n_t > expected.shape[0] or n_f != expected.shape[1] or n_a != expected.shape[2]:
            print(f"{path}: cube axes {data.shape[:3]} do not match the solutions {expected.shape[:3]}")
            status = 1
            continue
        sampled = np.asarray(data[:, :, :, args.plane][..., iy[inside], ix[inside]], np.float64)   # [t, f, a, d_in]
        want = expected[:n_t][..., inside]
        diff = sampled - want
        worst = np.unravel_index(np.argmax(np.abs(diff)), diff.shape)
        print(f"{os.path.basename(path)}: {n_t} x {n_f} x {n_a} cells, {n_y} x {n_x} pixels, plane {args.plane}; "
              f"{int(inside.sum())} of {len(inside)} patches inside the image")
        print(f"  screen - solution at the patch pixels: max |.| {np.abs(diff).max():.3e}, rms {np.sqrt((diff ** 2).mean()):.3e}, "
              f"worst at (time, freq, antenna, patch) = {tuple(int(w) for w in worst)}")
        if None not in (args.time, args.freq, args.antenna):
            t, f, a = args.time, args.freq, args.antenna
            print(f"  cell (time {t}, freq {f}, antenna {a}): patch, pixel (x, y), solution, screen, difference")
            names = [str(n) for n in np.asarray(sol.dir)[inside]]
            for k, name in enumerate(names):
                print(f"    {name:>16s} ({ix[inside][k]:5d}, {iy[inside][k]:5d})  {want[t, f, a, k]:+.6f}  "
                      f"{sampled[t, f, a, k]:+.6f}  {diff[t, f, a, k]:+.2e}")
        dump[os.path.basename(path)] = sampled
    if args.dump:
        np.savez_compressed(args.dump, **dump)
    return status


if __name__ == "__main__":
    sys.exit(main())

"""
TEST INFRASTRUCTURE (oracle): golden vectors for the gain ("amplitude + phase") path,
produced by the REFERENCE's own code under ``oracle.refshim``.

The repo's ``solutions.h5`` is phase-only, and there is no HDF5 writer in this image, so a
seeded synthetic ``amplitude000`` soltab (slow grid: 5 times x 3 freqs, XX/YY) is overlaid on
the real file in memory (``fake_h5parm.add_soltab``) and the reference is run with
``soltabname="gain000"`` (make_aterm_images.py:71-74) for both screen types.  The amplitude
inputs are stored in the fixture so the product reads exactly the same numbers.

    python -m oracle.make_golden --only gain
"""
import os
import shutil
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")
SUB_F = (0, 5, 11)
SUB_A = (0, 1, 30, 61)

TEST_ARGS = dict(
    bounds_deg=[124.565, 66.165, 127.895, 62.835],
    bounds_mid_deg=[126.23, 64.50],
    skymodel="skymodel.txt",
    solsetname="sol000",
    padding_fraction=0,
    cellsize_deg=0.2,
    ncpu=0,
)


def synthetic_amplitudes(times_ph, freqs_ph, n_ant, n_dir, seed=17):
    """Slow XX/YY amplitudes on a coarser grid than the phases: smooth over direction and
    station, a few flagged solutions and a few gross outliers (exercise _flag_outliers)."""
    rng = np.random.default_rng(seed)
    t_amp = np.linspace(times_ph[0] + 3.0, times_ph[-1] - 5.0, 5)
    f_amp = np.array([freqs_ph[1], freqs_ph[5], freqs_ph[10]]) + 1.0e3
    d_slope = rng.normal(0.0, 0.05, (n_dir, 2))
    a_gain = rng.normal(0.0, 0.03, (n_ant, 1, 2))
    val = (1.0 + a_gain[None, None] + d_slope[None, None, None]
           + 0.02 * np.sin(np.arange(5))[:, None, None, None, None]
           + 0.01 * np.arange(3)[None, :, None, None, None]
           + rng.normal(0.0, 0.01, (5, 3, n_ant, n_dir, 2)))
    weight = np.ones(val.shape, np.float32)
    weight[rng.random(val.shape) < 0.02] = 0.0
    out = rng.random(val.shape) < 0.004
    val[out] *= 3.0
    return t_amp, f_amp, val, weight


def golden_gain(mods):
    from oracle.refshim import fake_astropy, fake_h5parm

    tmp = tempfile.mkdtemp(prefix="ssf_golden_")
    for f in ("solutions.h5", "skymodel.txt"):
        shutil.copy(os.path.join(GOLDEN, f), tmp)
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        fake_h5parm.reset()
        h5 = fake_h5parm.H5parm("solutions.h5")
        ph = h5.get_solset("sol000").get_soltab("phase000")
        t_amp, f_amp, val, weight = synthetic_amplitudes(ph.time, ph.freq, len(ph.ant), len(ph.dir))
        fake_h5parm.add_soltab("solutions.h5", "sol000", "amplitude000", "amplitude",
                               ["time", "freq", "ant", "dir", "pol"],
                               [t_amp, f_amp, ph.ant, ph.dir, np.array(["XX", "YY"])], val, weight)
        out = dict(amp_time=t_amp, amp_freq=f_amp, amp_val=val, amp_weight=weight,
                   sub_f=SUB_F, sub_a=SUB_A)
        for screen_type in ("tessellated", "kl"):
            args = dict(TEST_ARGS, bounds_deg=list(TEST_ARGS["bounds_deg"]))
            mods["make_aterm_images"].make_aterm_image(
                "solutions.h5", soltabname="gain000", screen_type=screen_type, outroot=screen_type,
                smooth_deg=0.1 if screen_type == "tessellated" else 0.0, **args)
            cube = fake_astropy.open(f"{screen_type}_0.fits")[0].data
            out[f"cube_{screen_type}"] = cube[:, SUB_F][:, :, SUB_A].copy()
        args = dict(TEST_ARGS, bounds_deg=list(TEST_ARGS["bounds_deg"]))
        mods["make_aterm_images"].make_aterm_image(
            "solutions.h5", soltabname="gain000", screen_type="tessellated", outroot="plain",
            smooth_deg=0.0, **args)
        out["cube_tessellated_nosmooth"] = fake_astropy.open("plain_0.fits")[0].data[:, SUB_F][:, :, SUB_A].copy()
        store = fake_h5parm._CREATED["solutions.h5"]["sol000"]
        scr = store["amplitude_screen000"]._soltab
        res = store["amplitude_screen000resid"]._soltab
        out.update(amp_screen=scr._val, amp_screen_weights=scr._weight, amp_resid=res._val,
                   amp_order=res._weight[:, :, :, 0, :])
        np.savez_compressed(os.path.join(GOLDEN, "gain_c1.npz"), **out)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)
        fake_h5parm.reset()


def tec_case(seed=29, n_d=12, n_t=8, n_st=3):
    """Seeded dTEC-like values [station, dir, time]: a gradient over the pierce points per
    station plus noise, a few flagged solutions and gross outliers."""
    rng = np.random.default_rng(seed)
    dirs = np.column_stack([np.deg2rad(126.23 + rng.uniform(-2, 2, n_d)),
                            np.deg2rad(64.5 + rng.uniform(-1.5, 1.5, n_d))]).astype(np.float32)
    grad = rng.normal(0, 0.04, (n_st, 2))
    xy = np.rad2deg(dirs.astype(np.float64)) - [126.23, 64.5]
    val = (grad[:, None, None, 0] * xy[None, :, None, 0] + grad[:, None, None, 1] * xy[None, :, None, 1]
           + 0.01 * np.sin(np.arange(n_t))[None, None, :])
    sigma = np.array([0.004, 1.0, 1.6])[:n_st, None, None]     # reduced chi2 << 1, ~1, > 1: orders move
    val = val + sigma * rng.normal(0, 1.0, (n_st, n_d, n_t))
    weights = np.ones(val.shape, np.float32)
    weights[rng.random(val.shape) < 0.05] = 0.0
    out = rng.random(val.shape) < 0.02
    val[out] += 12.0 * np.broadcast_to(sigma, val.shape)[out]
    return dirs, val, weights


def golden_tec(mods):
    """_process_station(screen_type="tec") of the reference (stationscreen.py:549-562, 338-346,
    742-746) on the seeded case: the tec type is in the fit operator but no API path reaches it,
    so the operator itself is the pin."""
    ss = mods["stationscreen"]
    dirs, val, weights = tec_case()
    n_st, n_d, n_t = val.shape
    pp, _, _ = ss._calculate_piercepoints(np.zeros((1, 3)), dirs)
    full = ss._calculate_svd(pp, 100, 5.0 / 3.0, n_d)
    ss.FULL_MATRICES = full
    names = np.array([f"d{i}" for i in range(n_d)])
    out = dict(dirs=dirs, val=val, weights=weights, station_order=np.array([5.0, 6.0, 8.0]))
    res = [ss._process_station(val[a].copy(), pp, np.full(n_t, out["station_order"][a]), weights[a].copy(),
                               "tec", 2, 5.0, True, names, full, 5.0 / 3.0, 100) for a in range(n_st)]
    out.update(screen=np.stack([r[0] for r in res]), weights_out=np.stack([r[1] for r in res]),
               resid=np.stack([r[2] for r in res]), order=np.stack([r[3] for r in res]))
    np.savez_compressed(os.path.join(GOLDEN, "tec_station.npz"), **out)


GAP_AFTER = ((8, 200.0), (15, 500.0))       # (first slot after the gap, extra seconds)


def gapped_times(times):
    out = np.array(times, np.float64)
    for start, extra in GAP_AFTER:
        out[start:] += extra
    return out


def golden_gaps(mods):
    """Time gaps -> one FITS file per gap-free chunk (screen.py:283-294, 328-394): the repo's
    phase000 values on a time axis with two gaps (overlaid as phase001), tessellated screens,
    unsmoothed.  Stores per file the time-axis header cards and a sub-sampled cube, plus the
    file list."""
    from oracle.refshim import fake_astropy, fake_h5parm

    tmp = tempfile.mkdtemp(prefix="ssf_golden_")
    for f in ("solutions.h5", "skymodel.txt"):
        shutil.copy(os.path.join(GOLDEN, f), tmp)
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        fake_h5parm.reset()
        h5 = fake_h5parm.H5parm("solutions.h5")
        ph = h5.get_solset("sol000").get_soltab("phase000")
        times = gapped_times(ph.time)
        fake_h5parm.add_soltab("solutions.h5", "sol000", "phase001", "phase", ["time", "freq", "ant", "dir"],
                               [times, ph.freq, ph.ant, ph.dir], np.array(ph.val), np.array(ph.weight))
        args = dict(TEST_ARGS, bounds_deg=list(TEST_ARGS["bounds_deg"]))
        mods["make_aterm_images"].make_aterm_image(
            "solutions.h5", soltabname="phase001", screen_type="tessellated", outroot="gap", smooth_deg=0.0, **args)
        with open("gap.txt", encoding="utf8") as fh:
            names = [os.path.basename(ln.strip()) for ln in fh if ln.strip()]
        out = dict(times=times, files=np.array(names), sub_f=SUB_F, sub_a=SUB_A)
        for k, name in enumerate(names):
            hdu = fake_astropy.open(name)
            hdr = hdu[0].header
            out[f"naxis6_{k}"] = hdr["NAXIS6"]
            out[f"crval6_{k}"] = hdr["CRVAL6"]
            out[f"cdelt6_{k}"] = hdr["CDELT6"]
            out[f"cube_{k}"] = hdu[0].data[:, SUB_F][:, :, SUB_A].copy()
        np.savez_compressed(os.path.join(GOLDEN, "gaps_c1.npz"), **out)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)
        fake_h5parm.reset()


def golden_padding(mods):
    """Default padding_fraction (1.4) and a cell size that does not divide the width
    (make_aterm_images.py:88-97, :116-118; screen.py:180-181): image size, header and cube of
    the reference for the repo data, tessellated, unsmoothed, README-style bounds."""
    from oracle.refshim import fake_astropy, fake_h5parm

    tmp = tempfile.mkdtemp(prefix="ssf_golden_")
    for f in ("solutions.h5", "skymodel.txt"):
        shutil.copy(os.path.join(GOLDEN, f), tmp)
    cwd = os.getcwd()
    os.chdir(tmp)
    try:
        fake_h5parm.reset()
        mods["make_aterm_images"].make_aterm_image(
            "solutions.h5", soltabname="phase000", screen_type="tessellated", outroot="pad",
            bounds_deg="[127.895;62.835;124.565;66.165]", bounds_mid_deg="[126.23;64.50]",
            skymodel="skymodel.txt", solsetname="sol000", cellsize_deg=0.13, smooth_deg=0.0, ncpu=0)
        hdu = fake_astropy.open("pad_0.fits")
        hdr = hdu[0].header
        out = dict(sub_f=SUB_F, sub_a=SUB_A, cube=hdu[0].data[:, SUB_F][:, :, SUB_A].copy(),
                   header_keys=np.array([k for k, _ in hdr.items()]),
                   header_vals=np.array([str(v) for _, v in hdr.items()]),
                   template=fake_astropy.open("pad_template.fits")[0].data[0, 0, 0, 0].copy())
        np.savez_compressed(os.path.join(GOLDEN, "padding_c1.npz"), **out)
    finally:
        os.chdir(cwd)
        shutil.rmtree(tmp)
        fake_h5parm.reset()


JOBS = {"gain": golden_gain, "tec": golden_tec, "gaps": golden_gaps, "padding": golden_padding}


# ------------------------------------------------------------------ phase fit at real direction counts
PHASE_SHAPES = {"phase45": dict(n_dirs=45, seed=45), "phase100": dict(n_dirs=100, seed=100)}
PIX_SAMPLE = tuple(range(5, 512, 67))        # 8 rows x 8 columns of the 512 x 512 grid


def phase_case(n_dirs, seed, n_times=4, n_freqs=2, n_stations=6):
    """Seeded LOFAR-like case at the direction counts of BASELINE.json's configs (45 / 100) on the
    512 x 512 grid of configs[2]: 3 % of the solutions flagged (weight 0), 1 % gross phase
    outliers (+2.5 rad) for the 5-sigma circular clip -- which can only fire above ~30 directions
    -- and station noise levels from far below to far above the level at which the reduced
    chi-square crosses 1, so the screen orders move in both directions."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import synthetic_cases

    case = synthetic_cases.make_case("lofar", seed=seed, n_times=n_times, n_freqs=n_freqs,
                                     n_stations=n_stations, n_dirs=n_dirs, flag_fraction=0.03)
    rng = np.random.default_rng(seed + 1)
    sol = case.solutions
    noise = np.array([0.0, 0.004, 0.05, 0.3, 1.0, 1.6])[:n_stations]
    sol.val = sol.val + noise[None, None, :, None] * rng.normal(0, 1.0, sol.val.shape)
    sol.val[rng.random(sol.val.shape) < 0.01] += 2.5
    sol.weight[:, :, 0, :] = 1.0                  # station 0 stays the reference station
    sol.weight[:, 1, 3, :] = 0.0                  # a fully flagged (freq, station): skipped by the fit
    # The circular chi-square of the phase type is (1 - hypot(<sin^2>, <cos^2>)) * sum(w)
    # (stationscreen.py:353-387): at most 0.29 per unit of weight, so with unit weights the
    # reduced chi-square (:729-731) stays below 1 and the order never leaves its starting value.
    # It scales with the weights, though, and h5parm weights need not be 1: station 3 carries
    # inverse-variance-like weights of 25, station 4 weights drawn from {1, 4, 9} (non-uniform:
    # the general normal-equation path), and the two noisy stations keep only 23 / 24 directions
    # in two time slots (the order is capped at n_unflagged - 1 there).
    if n_stations > 4:
        sol.weight[:, :, 0, :] = 30.0             # ... and station 0 still has the largest weight sum
        sol.weight[:, :, 3, :] *= 25.0
        sol.weight[:, :, 4, :] *= rng.choice([1.0, 4.0, 9.0], size=sol.weight[:, :, 4, :].shape).astype(np.float32)
    for a in range(max(0, n_stations - 2), n_stations):
        for t, keep in ((1, 23), (2, 24)):
            drop = rng.permutation(n_dirs)[keep:]
            sol.weight[t, :, a, drop] = 0.0
    return case


def golden_phase(mods, tag):
    """The reference's make_aterm_image(screen_type="kl") on ``phase_case``: fit soltabs
    (screen, weights after flagging, residuals, orders) and the cube at an 8 x 8 pixel sample."""
    from oracle import refrun
    from oracle.refshim import fake_astropy, fake_h5parm

    case = phase_case(**PHASE_SHAPES[tag])
    sol = case.solutions
    tmp = tempfile.mkdtemp(prefix="ssf_golden_")
    try:
        fake_h5parm.reset()
        files, tabs = refrun.run_case(case, "kl", tmp, ncpu=0, name=tag)
        cube = fake_astropy.open(files[0])[0].data
        scr, res = tabs["phase_screen000"], tabs["phase_screen000resid"]
        assert cube.shape[-1] == 512 and len(files) == 1
        ix = np.array(PIX_SAMPLE)
        out = dict(val=np.asarray(sol.val), weight=np.asarray(sol.weight), time=sol.time, freq=sol.freq,
                   ant=np.array(sol.ant), dir=np.array(sol.dir),
                   source_rad=np.array(sol.source_positions), station_xyz=np.array(sol.station_positions),
                   patch_radec=case.patch_radec, rad=case.rad, dec=case.dec, width_deg=case.width_deg,
                   cellsize_deg=case.cellsize_deg, pix=ix,
                   screen=scr._val, screen_weights=scr._weight, resid=res._val, order=res._weight[..., 0],
                   piercepoints=np.array(scr.obj.piercepoint),
                   midra=scr.obj._v_attrs["midra"], middec=scr.obj._v_attrs["middec"],
                   cube_sample=cube[..., ix[:, None], ix[None, :]].copy())
        flagged = (out["screen_weights"] != out["weight"]).sum()
        moved = (out["order"] != out["order"][0:1]).sum()
        print(f"  {tag}: {flagged} solutions flagged by the clip, {moved} systems with a moved order, "
              f"orders seen: {np.unique(out['order'])}")
        assert flagged > 0, "the case must trigger the 5-sigma clip"
        assert moved > 0, "the case must move the screen orders"
        np.savez_compressed(os.path.join(GOLDEN, f"{tag}.npz"), **out)
    finally:
        shutil.rmtree(tmp)
        fake_h5parm.reset()


JOBS.update({"phase45": lambda mods: golden_phase(mods, "phase45"),
             "phase100": lambda mods: golden_phase(mods, "phase100")})


def tec_run_case(seed=31, n_t=8, n_f=2, n_a=4, n_d=12):
    """A dTEC soltab [time, freq, ant, dir]: ``tec_case`` per frequency (station 0 is the quiet
    reference), stations a few km apart."""
    rng = np.random.default_rng(seed)
    dirs, val, weights = tec_case(seed=seed, n_d=n_d, n_t=n_t, n_st=n_a - 1)
    val_t = np.zeros((n_t, n_f, n_a, n_d))
    wt = np.ones((n_t, n_f, n_a, n_d), np.float32)
    for f in range(n_f):
        val_t[:, f, 1:, :] = val.transpose(2, 0, 1) * (1.0 + 0.1 * f)
        wt[:, f, 1:, :] = weights.transpose(2, 0, 1)
    val_t[:, :, 0, :] = rng.normal(0, 0.001, (n_t, n_f, n_d))
    pos = (np.array([3826896.2, 460979.5, 5064658.0]) + rng.normal(0, 5e3, (n_a, 3))).astype(np.float32)
    return dirs, val_t, wt, pos


def golden_tec_run(mods):
    """stationscreen.run on a soltab of TYPE "tec" (the whole operation: referencing, per-station
    orders, _process_single_freq, result soltabs) and the screen images K . white of
    calculate_kl_screen (kl_screen.py:411-449) on a 16 x 16 grid: what an aterm_type="tec" cube
    (processing_utils.py:187-190) of those screens holds.  No API path of the reference reaches
    either with a tec soltab, so the operators themselves are the pin."""
    from types import SimpleNamespace

    from oracle import refrun
    from oracle.port import wcs as owcs
    from oracle.refshim import fake_h5parm

    ss, kls, misc = mods["stationscreen"], mods["kl_screen"], mods["processing_utils"]
    dirs, val, wt, pos = tec_run_case()
    n_t, n_f, n_a, n_d = val.shape
    ants, names = [f"ST{i:03d}" for i in range(n_a)], [f"[Patch_{i}]" for i in range(n_d)]
    sol = SimpleNamespace(val=val, weight=wt, time=4.9e9 + 8.0 * np.arange(n_t), freq=1.2e8 + 4e6 * np.arange(n_f),
                          ant=ants, dir=names, station_dict=dict(zip(ants, pos)), source_dict=dict(zip(names, dirs)))
    fake_h5parm.reset()
    try:
        refrun.register(sol, "tec.h5", soltab_name="tec000", soltype="tec")
        solset = fake_h5parm.H5parm("tec.h5").get_solset("sol000")
        soltab = solset.get_soltab("tec000")
        ref_ind = misc.get_reference_station(soltab, 10)
        ss.run(soltab, "tec_screen000", order=min(20, n_d - 1), ref_ant=ref_ind, scale_order=True,
               adjust_order=True, ncpu=1)
        store = fake_h5parm._CREATED["tec.h5"]["sol000"]
        scr, res = store["tec_screen000"]._soltab, store["tec_screen000resid"]._soltab
        pp = np.array(scr.obj.piercepoint)
        rad, dec, n_pix, cell = 126.23, 64.5, 16, 0.25
        x, y = owcs.kl_grid_coords(rad, dec, (n_pix - 0.5) * cell, cell, scr.obj._v_attrs["midra"],
                                   scr.obj._v_attrs["middec"])
        images = np.zeros((n_t, n_f, n_a, n_pix, n_pix))
        kls.PIERCEPOINTS, kls.X_COORD, kls.Y_COORD = pp, x, y
        for f in range(n_f):
            for a in range(n_a):
                kls.SCREEN_PH = scr._val[:, f, a, :].T.copy()                    # [dir, time], kl_screen.py:263-271
                shared = np.zeros((n_pix, n_pix, n_t))
                kls.VAR_DICT = {"shared_val": shared, "val_shape": shared.shape}     # as make_matrix sets it (:293)
                for k in range(n_t):
                    kls.calculate_kl_screen(k, n_d, 5.0 / 3.0, 100, "ph")
                images[:, f, a] = shared.transpose(2, 1, 0)                     # [t, y, x] (the .T of :368-373)
        out = dict(dirs=dirs, val=val, weight=wt, station_xyz=pos, time=sol.time, freq=sol.freq,
                   ref_ind=ref_ind, screen=scr._val, screen_weights=scr._weight, resid=res._val,
                   order=res._weight[..., 0], piercepoints=pp, rad=rad, dec=dec, n_pix=n_pix, cellsize_deg=cell,
                   images=images)
        assert (out["screen_weights"] != wt).any() and len(np.unique(out["order"])) > 2
        np.savez_compressed(os.path.join(GOLDEN, "tec_run.npz"), **out)
    finally:
        fake_h5parm.reset()


JOBS["tec_run"] = golden_tec_run

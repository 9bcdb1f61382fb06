"""
Parity checks shared by the emulated (``not gpu``) and the real (``gpu``) test modules: each
takes an ``Engine`` and compares the C-ABI kernels with ``oracle.port`` on seeded inputs.

Tolerances (stated once, used everywhere):
  * template indices, Voronoi fill, NaN fix, byte swap, flags, orders: BIT-EXACT
  * KL fit (FP64): 1e-9 absolute on white / residual (rad)
  * KL screens: 1e-5 absolute on the phase-derived planes cos/sin (north_star: 1e-5 on phase)
  * Gaussian smoothing: 1e-6 absolute (float32 accumulation vs scipy's float64 line buffer)
"""
import os

import numpy as np
import torch

from oracle.port import klfit, screens, wcs as owcs

TOL_FIT = 1e-9
TOL_KL = 1e-5
TOL_SMOOTH = 1e-6


def expect_formats(native, run, planes2=True):
    """The fused output formats of a cube kernel (SSF_OUT_BIG_ENDIAN, SSF_OUT_PLANES2) against
    its native 4-plane little-endian output: same values, bit for bit, only stored differently.
    ``run(out_flags, planes)`` returns the slab as a numpy float32 array [ncol, planes, ...]."""
    for flags in (1, 2, 3) if planes2 else (1,):
        planes = 2 if flags & 2 else 4
        got = run(flags, planes)
        want = native[:, :planes]
        if flags & 1:
            got = got.view(np.uint32).byteswap().view(np.float32)
        assert np.array_equal(got.view(np.uint32), np.ascontiguousarray(want).view(np.uint32)), flags


def make_directions(rng, n_d, spread=2.0):
    return np.column_stack([np.deg2rad(126.23 + rng.uniform(-spread, spread, n_d) / np.cos(np.deg2rad(64.5))),
                            np.deg2rad(64.5 + rng.uniform(-spread, spread, n_d))]).astype(np.float32)


def mask_of(bits):
    m = sum(1 << int(b) for b in bits)
    lo, hi = m & (2 ** 64 - 1), m >> 64
    return [lo - 2 ** 64 if lo >= 2 ** 63 else lo, hi - 2 ** 64 if hi >= 2 ** 63 else hi]


def check_decompose(eng, n_d=9, seed=5):
    rng = np.random.default_rng(seed)
    pp, _, _ = klfit.piercepoints(make_directions(rng, n_d))
    subsets = [list(range(n_d)), [i for i in range(n_d) if i not in (1, 4)], [0, 2, n_d - 1]]
    masks = torch.tensor([mask_of(s) for s in subsets], dtype=torch.int64, device=eng.device)
    c, p, u, sv = eng.kl_decompose(eng.tensor(pp[:, :2], torch.float64), masks)
    for k, sub in enumerate(subsets):
        n = len(sub)
        c_o, p_o, u_o = klfit.calculate_svd(pp[sub])
        sv_o = np.linalg.svd(c_o, compute_uv=False)
        # pow() differs by an ulp or two between libm, numpy's SIMD loops and CUDA
        assert np.allclose(c[k, :n, :n].cpu().numpy(), c_o, rtol=1e-13, atol=0)
        assert np.allclose(sv[k, :n].cpu().numpy(), sv_o, rtol=1e-12, atol=1e-10)
        scale = np.abs(p_o).max()
        assert np.abs(p[k, :n, :n].cpu().numpy() - p_o).max() < 1e-10 * max(scale, 1.0)
        u_k = u[k, :n, :n].cpu().numpy()
        signs = np.sign(np.sum(u_k * u_o, axis=0))
        assert np.abs(u_k * signs - u_o).max() < 1e-9
        assert np.all(c[k, n:, :].cpu().numpy() == 0) and np.all(p[k, :, n:].cpu().numpy() == 0)


def check_pattern_cache_decompose(eng, n_d=11, seed=6, max_flagged=3, n_sys=6):
    """Decompositions handed out by the device pattern cache (warm-started from slot 0's
    eigenvectors) against the reference's SVD / pinv of the same subsets."""
    from ska_sdp_screen_fitting_b200.engine import PatternCache
    rng = np.random.default_rng(seed)
    pp, _, _ = klfit.piercepoints(make_directions(rng, n_d))
    cache = PatternCache(eng, eng.tensor(pp[:, :2], torch.float64), max_bytes=64 << 20)
    subsets = [list(range(n_d))]
    for s in range(n_sys - 1):
        drop = rng.choice(n_d, size=1 + s % max_flagged, replace=False)
        subsets.append([i for i in range(n_d) if i not in drop])
    subsets.append([0, n_d // 2, n_d - 1])                      # heavily flagged
    masks = torch.tensor([mask_of(s) for s in subsets], dtype=torch.int64, device=eng.device)
    pat = cache.lookup(masks).cpu().numpy()
    eng.check_fit_status()
    assert pat[0] == 0 and len(set(pat.tolist())) == len({tuple(s) for s in subsets})
    for k, sub in enumerate(subsets):
        n = len(sub)
        c_o, p_o, u_o = klfit.calculate_svd(pp[sub])
        slot = int(pat[k])
        assert np.allclose(cache.c[slot, :n, :n].cpu().numpy(), c_o, rtol=1e-13, atol=0)
        sv_o = np.linalg.svd(c_o, compute_uv=False)
        assert np.allclose(cache.sv[slot, :n].cpu().numpy(), sv_o, rtol=1e-12, atol=1e-10)
        scale = np.abs(p_o).max()
        assert np.abs(cache.p[slot, :n, :n].cpu().numpy() - p_o).max() < 1e-10 * max(scale, 1.0)
        u_k = cache.u[slot, :n, :n].cpu().numpy()
        signs = np.sign(np.sum(u_k * u_o, axis=0))
        assert np.abs(u_k * signs - u_o).max() < 1e-9


def check_pattern_cache_pressure(eng, n_d=8, seed=41):
    """A pass with far more flagged systems than cache slots: 50 systems that share
    ONE flag pattern fit in a 4-slot cache, and 12 distinct patterns go through it in runs of
    max_slots - 1 systems (the cache forgets between runs) -- both equal a roomy cache."""
    from ska_sdp_screen_fitting_b200.engine import PatternCache
    rng = np.random.default_rng(seed)
    pp, _, _ = klfit.piercepoints(make_directions(rng, n_d))
    ppt = eng.tensor(pp[:, :2], torch.float64)
    per = (3 * n_d * n_d + n_d) * 8 + 16
    for n_sys, distinct in ((50, 1), (36, 12)):
        val = rng.normal(0, 0.3, (n_sys, n_d))
        weight = np.ones((n_sys, n_d), np.float32)
        for i in range(n_sys):
            k = i % distinct
            weight[i, k % n_d] = 0.0
            if k >= n_d:
                weight[i, (k + 3) % n_d] = 0.0
        so = np.full(n_sys, 4.0)
        outs = []
        for max_bytes in (4 * per, 64 << 20):
            cache = PatternCache(eng, ppt, max_bytes=max_bytes)
            assert (cache.max_slots == 4) == (max_bytes == 4 * per)
            w = eng.tensor(weight, torch.float32)
            white, resid, order = eng.kl_fit(eng.tensor(val, torch.float64), w, eng.tensor(so, torch.float64),
                                             ppt, cache=cache)
            outs.append((white.cpu().numpy(), resid.cpu().numpy(), order.cpu().numpy()))
            assert cache.stats()[1] == 0
        for a, b in zip(*outs):
            assert np.array_equal(a, b)
        assert np.abs(outs[0][0]).max() > 0
    # a lookup that really holds more patterns than slots: reported, systems without a slot get -1
    cache = PatternCache(eng, ppt, max_bytes=4 * per)
    masks = torch.tensor([mask_of([d for d in range(n_d) if d != k]) for k in range(6)], dtype=torch.int64,
                         device=eng.device)
    pat = cache.lookup(masks).cpu().numpy()
    assert cache.stats()[1] == 1 and (pat == -1).sum() == 3 and sorted(pat[pat > 0]) == [1, 2, 3]
    import pytest
    from ska_sdp_screen_fitting_b200._lib import SsfError
    with pytest.raises(SsfError):
        eng.check_fit_status()
    eng.check_fit_status()                                   # the flag is cleared by the report


def synthetic_fit_case(seed, n_t, n_f, n_a, n_d, flag_fraction=0.03):
    rng = np.random.default_rng(seed)
    dirs = make_directions(rng, n_d)
    pos = (np.array([3826896.2, 460979.5, 5064658.0]) + rng.normal(0, 2e4, (n_a, 3))).astype(np.float32)
    pp, _, _ = klfit.piercepoints(dirs)
    grad = rng.normal(0, 3e-4, (n_a, 2))
    val = (grad[None, None, :, 0, None] * pp[None, None, None, :, 0]
           + grad[None, None, :, 1, None] * pp[None, None, None, :, 1]
           + rng.normal(0, 0.02, (n_t, n_f, n_a, n_d)))
    val[rng.random(val.shape) < 0.01] += 2.5                 # outliers for the 5-sigma clip
    weight = np.ones(val.shape, np.float32)
    weight[rng.random(val.shape) < flag_fraction] = 0.0
    weight[:, :, 0, :] = 1.0                                 # keep station 0 the reference
    if n_a > 3:
        weight[:, 0, 3, :] = 0.0                             # a fully flagged (freq, station)
    return val, weight, dirs, pos


def run_engine_fit(eng, val, weight, dirs, pos):
    """The KLScreen.fit recipe on raw arrays (kept here so the check does not depend on files)."""
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

    n_t, n_f, n_a, n_d = val.shape
    sol = Solutions.from_arrays(val, weight, np.arange(n_t) * 8.0, 1e8 + np.arange(n_f) * 1e6,
                                [f"a{i}" for i in range(n_a)], [f"d{i}" for i in range(n_d)], dirs, pos)
    scr = KLScreen("t", None, None, 126.23, 64.5, 1.0, 1.0)
    scr.solutions = sol
    scr._engine = eng
    scr.fit()
    return scr


def check_fit(eng, seed=7, n_t=3, n_f=2, n_a=5, n_d=40):
    # a 5-sigma clip over directions can only trigger with >~30 directions: (n-1)/sqrt(n) > 5
    val, weight, dirs, pos = synthetic_fit_case(seed, n_t, n_f, n_a, n_d)
    scr = run_engine_fit(eng, val, weight, dirs, pos)
    ref = klfit.get_reference_station(weight, 10)
    out = klfit.run_phase(val, weight.copy(), dirs, pos, ref, min(20, n_d - 1))
    assert np.abs(scr.piercepoints - out["pp"]).max() < 1e-9
    assert np.abs(scr.vals_ph - out["screen"]).max() < TOL_FIT
    assert np.abs(scr.residuals - out["resid"]).max() < TOL_FIT
    assert np.array_equal(scr.screen_order, out["order"])
    assert np.array_equal(scr.weights_out, out["weights"])
    assert (out["weights"] != weight).any(), "the case must exercise outlier flagging"
    return scr, out


def check_fit_small_weights(eng, seed=41, n_t=3, n_f=2, n_a=4, n_d=24):
    """Weights in (0, 1e-3]: U_k^T W U_k can have eigenvalues at or below the cutoff of
    pinv(..., rcond=1e-3) (stationscreen.py:501-504), which then drops them.  The kernel takes
    its general (one-sided Jacobi) path for such systems; against the port's scipy pinv."""
    val, weight, dirs, pos = synthetic_fit_case(seed, n_t, n_f, n_a, n_d, flag_fraction=0.02)
    rng = np.random.default_rng(seed + 1)
    weight = weight.astype(np.float64)
    weight[weight > 0] = rng.uniform(0.5, 2.0, int((weight > 0).sum()))
    # station 1: a handful of tiny weights (all eigenvalues stay above the cutoff);
    # station 2: all but 5 directions tiny, order capped at 20 > 5 -> eigenvalues below the
    # cutoff are dropped; one system with every weight tiny -> pinv = 0 -> zero screen
    tiny = rng.random((n_t, n_f, n_d)) < 0.2
    weight[:, :, 1, :] = np.where(tiny & (weight[:, :, 1, :] > 0), 3e-4, weight[:, :, 1, :])
    keep = rng.permutation(n_d)[:5]
    w2 = np.full((n_t, n_f, n_d), 5e-4)
    w2[:, :, keep] = 1.0
    weight[:, :, 2, :] = np.where(weight[:, :, 2, :] > 0, w2, 0.0)
    weight[0, 0, 2, :] = 2e-4
    weight = weight.astype(np.float32)
    scr = run_engine_fit(eng, val, weight, dirs, pos)
    ref = klfit.get_reference_station(weight, 10)
    out = klfit.run_phase(val, weight.copy(), dirs, pos, ref, min(20, n_d - 1))
    assert np.array_equal(scr.screen_order, out["order"])
    assert np.array_equal(scr.weights_out, out["weights"])
    assert np.abs(scr.vals_ph - out["screen"]).max() < TOL_FIT * 10
    assert np.abs(scr.residuals - out["resid"]).max() < TOL_FIT * 10
    return scr, out


def check_fit_amplitude(eng, seed=23, n_t=6, n_f=2, n_a=4, n_d=24):
    """Amplitude-type fit (log10 screens, per-station rms clip over all times, niter=3)
    against the port of stationscreen.run / _process_station / _fit_screen."""
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen, amplitude_order

    rng = np.random.default_rng(seed)
    dirs = make_directions(rng, n_d)
    pos = (np.array([3826896.2, 460979.5, 5064658.0]) + rng.normal(0, 2e4, (n_a, 3))).astype(np.float32)
    pp, _, _ = klfit.piercepoints(dirs)
    grad = rng.normal(0, 2e-4, (n_a, 2, 2))
    log_amp = (grad[None, None, :, None, 0, :] * pp[None, None, None, :, 0, None]
               + grad[None, None, :, None, 1, :] * pp[None, None, None, :, 1, None]
               + rng.normal(0, 0.004, (n_t, n_f, n_a, n_d, 2)))
    val = 10 ** log_amp
    out_mask = rng.random(val.shape) < 0.004
    val[out_mask] *= 1.6                                     # gross outliers for the 5-sigma clip
    weight = np.ones(val.shape, np.float32)
    weight[rng.random(val.shape) < 0.03] = 0.0
    weight[:, 1, 2, :, 0] = 0.0                              # a fully flagged (freq, station, pol)
    sol = Solutions.from_arrays(np.zeros((n_t, n_f, n_a, n_d)), np.ones((n_t, n_f, n_a, n_d), np.float32),
                                np.arange(n_t) * 8.0, 1e8 + np.arange(n_f) * 1e6,
                                [f"a{i}" for i in range(n_a)], [f"d{i}" for i in range(n_d)], dirs, pos)
    scr = KLScreen("t", None, None, 126.23, 64.5, 1.0, 1.0)
    scr._engine = eng
    scr.setup(sol)
    white, resid, order, w_out = scr.fit_amplitude_device(eng.tensor(val, torch.float64),
                                                          eng.tensor(weight, torch.float32))
    want = klfit.run_amplitude(val, weight.copy(), dirs, amplitude_order(n_d))
    assert np.array_equal(w_out.cpu().numpy(), want["weights"])
    assert (want["weights"] != weight).any(), "the case must exercise outlier flagging"
    assert np.array_equal(order.cpu().numpy(), want["order"])
    assert np.abs(white.cpu().numpy() - want["screen"]).max() < TOL_FIT
    # the residuals are amplitudes (not logs) and scale with them: relative above 1
    assert (np.abs(resid.cpu().numpy() - want["resid"]) <= TOL_FIT * np.maximum(1.0, np.abs(want["resid"]))).all()
    return want


def check_fit_tec_golden(eng, golden_dir):
    """tec-type fit (values fitted as they are, rms clip on the plain residual, plain chi2)
    against the outputs of the reference's own _process_station(screen_type="tec")
    (tests/golden/tec_station.npz, oracle/make_golden_synth.py::golden_tec)."""
    g = np.load(os.path.join(golden_dir, "tec_station.npz"))
    n_st, n_d, n_t = g["val"].shape
    pp, _, _ = klfit.piercepoints(g["dirs"])
    val = eng.tensor(np.ascontiguousarray(g["val"].transpose(2, 0, 1)).reshape(n_t * n_st, n_d), torch.float64)
    wts = eng.tensor(np.ascontiguousarray(g["weights"].transpose(2, 0, 1)).reshape(n_t * n_st, n_d), torch.float32)
    st_order = eng.tensor(np.tile(g["station_order"], n_t), torch.float64)
    white, resid, order = eng.kl_fit(val, wts, st_order, eng.tensor(pp[:, :2], torch.float64), niter=2,
                                     screen_type="tec", ngroups=n_st)
    back = lambda t, shape: t.cpu().numpy().reshape(shape)
    assert np.array_equal(back(wts, (n_t, n_st, n_d)).transpose(1, 2, 0), g["weights_out"])
    assert (g["weights_out"] != g["weights"]).any(), "the case must exercise outlier flagging"
    assert np.array_equal(back(order, (n_t, n_st)).T, g["order"])
    assert (g["order"] != g["station_order"][:, None]).any(), "the case must exercise the order state machine"
    assert np.abs(back(white, (n_t, n_st, n_d)).transpose(1, 2, 0) - g["screen"]).max() < TOL_FIT
    assert np.abs(back(resid, (n_t, n_st, n_d)).transpose(1, 2, 0) - g["resid"]).max() < TOL_FIT * 20


def check_kl_gain_eval(eng, n=24, n_d=9, ncol=7, seed=31, cell=0.1, variant=0):
    """ssf_kl_eval_gain against the float64 oracle: 10^Axx / 10^Ayy times cos / sin."""
    rng = np.random.default_rng(seed)
    width = (n - 0.5) * cell
    pp, mid_ra, mid_dec = klfit.piercepoints(make_directions(rng, n_d, spread=0.4 * width))
    x, y = owcs.kl_grid_coords(126.23, 64.5, width, cell, mid_ra, mid_dec)
    c_o, p_o, _ = klfit.calculate_svd(pp)
    white = [rng.normal(0, s, (ncol, n_d)) @ p_o.T for s in (0.7, 0.05, 0.05)]
    ppt = eng.tensor(pp[:, :2], torch.float64)
    full = torch.tensor([mask_of(range(n_d))], dtype=torch.int64, device=eng.device)
    c, p, _, _ = eng.kl_decompose(ppt, full)
    bt = eng.kl_basis(eng.tensor(x, torch.float64), eng.tensor(y, torch.float64), ppt, p[0].contiguous())
    stacked = np.stack(white, axis=1).reshape(-1, n_d)
    g3 = eng.kl_coeffs(eng.tensor(stacked, torch.float64), c[0].contiguous()).view(ncol, 3, -1)
    slab = eng.empty((ncol, 4, n, n), torch.float32)
    eng.kl_eval_gain(bt, g3, n_d, slab, variant)
    want = screens.kl_make_matrix(x, y, pp, white[0], white[1], white[2])
    got = slab.cpu().numpy()
    assert (np.abs(got - want) <= TOL_KL * np.maximum(1.0, np.abs(want))).all()
    expect_formats(got, lambda flags, planes: eng.kl_eval_gain(
        bt, g3, n_d, eng.empty((ncol, planes, n, n), torch.float32), variant, flags).cpu().numpy(), planes2=False)


def check_kl_eval(eng, n=24, n_d=9, ncol=11, seed=3, variant=0, cell=0.1):
    """Directions are drawn in the central 80 % of the field (the geometry of BASELINE.json's
    synthetic configs): the screen is an interpolant there, not an extrapolant."""
    rng = np.random.default_rng(seed)
    width = (n - 0.5) * cell
    pp, mid_ra, mid_dec = klfit.piercepoints(make_directions(rng, n_d, spread=0.4 * width))
    x, y = owcs.kl_grid_coords(126.23, 64.5, width, cell, mid_ra, mid_dec)
    assert len(x) == n
    c_o, p_o, _ = klfit.calculate_svd(pp)
    fit = rng.normal(0, 0.7, (ncol, n_d))
    white = fit @ p_o.T                                      # white = pinvC fit, as the fit produces
    ppt = eng.tensor(pp[:, :2], torch.float64)
    full = torch.tensor([mask_of(range(n_d))], dtype=torch.int64, device=eng.device)
    c, p, _, _ = eng.kl_decompose(ppt, full)
    bt = eng.kl_basis(eng.tensor(x, torch.float64), eng.tensor(y, torch.float64), ppt, p[0].contiguous())
    g = eng.kl_coeffs(eng.tensor(white, torch.float64), c[0].contiguous())
    slab = eng.empty((ncol, 4, n, n), torch.float32)
    eng.kl_eval(bt, g, n_d, slab, variant)
    want = screens.kl_make_matrix(x, y, pp, white)
    got = slab.cpu().numpy()
    assert np.abs(got - want).max() < TOL_KL
    assert np.array_equal(got[:, 0], got[:, 2]) and np.array_equal(got[:, 1], got[:, 3])
    expect_formats(got, lambda flags, planes: eng.kl_eval(
        bt, g, n_d, eng.empty((ncol, planes, n, n), torch.float32), variant, flags).cpu().numpy())
    # dTEC cube: the screen values themselves (no cos/sin), one plane per column
    phi = screens.kl_screen_values(x, y, pp, white)
    for flags in (0, 1):
        tec = eng.kl_eval_tec(bt, g, n_d, eng.empty((ncol, n, n), torch.float32), variant, flags).cpu().numpy()
        if flags:
            tec = tec.view(np.uint32).byteswap().view(np.float32)
        assert np.abs(tec - phi).max() < TOL_KL * max(1.0, np.abs(phi).max())
    return float(np.abs(got - want).max())


def check_voronoi(eng, n=20, n_d=7, ncol=13, seed=9):
    rng = np.random.default_rng(seed)
    px = rng.uniform(-2, n + 2, n_d)
    py = rng.uniform(-2, n + 2, n_d)
    px[1], py[1] = 4.0, 6.0                                  # exact ties along x = 5 with site 2
    px[2], py[2] = 6.0, 6.0
    tmpl = eng.voronoi_assign(eng.tensor(px, torch.float64), eng.tensor(py, torch.float64), n, n)
    want_t = screens.nearest_patch_template(px, py, n)
    assert np.array_equal(tmpl.cpu().numpy(), want_t)
    phases = rng.normal(0, 1.0, (ncol, n_d))
    phases[ncol - 1, n_d - 1] = np.nan
    lut = eng.voronoi_lut(eng.tensor(phases, torch.float64))
    for fix in (False, True):
        slab = eng.empty((ncol, 4, n, n), torch.float32)
        eng.voronoi_fill(tmpl, lut, slab, fix)
        want = screens.voronoi_make_matrix(want_t, phases).astype(np.float32)
        if fix:
            want = screens.nan_fix(want)
        assert np.array_equal(slab.cpu().numpy(), want, equal_nan=True)
        expect_formats(slab.cpu().numpy(), lambda flags, planes: eng.voronoi_fill(
            tmpl, lut, eng.empty((ncol, planes, n, n), torch.float32), fix, flags).cpu().numpy())
    return tmpl, lut, want_t, phases


def check_voronoi_gain(eng, n=20, n_d=7, ncol=13, seed=19):
    """ssf_voronoi_lut_gain + ssf_voronoi_fill_gain: bit-exact against the oracle."""
    rng = np.random.default_rng(seed)
    want_t = screens.nearest_patch_template(rng.uniform(0, n, n_d), rng.uniform(0, n, n_d), n)
    tmpl = eng.tensor(want_t, torch.int32)
    phases = rng.normal(0, 1.0, (ncol, n_d))
    amp = rng.uniform(0.5, 1.5, (ncol, n_d, 2))
    phases[ncol - 1, n_d - 1] = np.nan
    amp[0, 0, 1] = np.nan
    lut4 = eng.voronoi_lut_gain(eng.tensor(phases, torch.float64), eng.tensor(amp, torch.float64))
    for fix in (False, True):
        slab = eng.empty((ncol, 4, n, n), torch.float32)
        eng.voronoi_fill_gain(tmpl, lut4, slab, fix)
        want = screens.voronoi_make_matrix(want_t, phases, amp).astype(np.float32)
        if fix:
            want = screens.nan_fix(want)
        assert np.array_equal(slab.cpu().numpy(), want, equal_nan=True)
        expect_formats(slab.cpu().numpy(), lambda flags, planes: eng.voronoi_fill_gain(
            tmpl, lut4, eng.empty((ncol, planes, n, n), torch.float32), fix, flags).cpu().numpy(), planes2=False)


def check_smoothed_voronoi_gain(eng, n=24, n_d=6, n_t=3, sigma=1.5, seed=23, variant=0):
    """Smoothed gain Voronoi through linearity (a gain slab seen as 2 x ncol plane pairs of the
    phase-only contraction, VoronoiScreen._fill_columns) against the reference's order of
    operations -- fill, gaussian_filter per plane, NaN fix (voronoi_screen.py:177-214,
    screen.py:353-378) -- from the float64 oracle, and against the direct device path."""
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    rng = np.random.default_rng(seed)
    want_t = screens.nearest_patch_template(rng.uniform(0, n, n_d), rng.uniform(0, n, n_d), n)
    ncol = n_t * 2
    phases = rng.normal(0, 1.0, (ncol, n_d))
    amp = rng.uniform(0.5, 1.5, (ncol, n_d, 2))
    want = screens.voronoi_make_matrix(want_t, phases, amp).astype(np.float32)
    want = screens.nan_fix(np.stack([screens.smooth_block(want[c], sigma) for c in range(ncol)]))

    def screen(mode):
        scr = VoronoiScreen("g", None, None, 0.0, 0.0, 1.0, 1.0, amplitude_soltab_name="amplitude000")
        scr._engine = eng
        scr.smooth_mode = mode
        scr._tmpl_dev = eng.tensor(want_t, torch.int32)
        scr._tmpl_key = "fixed"
        scr._prepare = lambda *a: None
        scr.source_names = [f"p{i}" for i in range(n_d)]
        scr._has_nan = False
        scr._dev_amp = eng.tensor(amp, torch.float64)
        return scr

    cols, amps = eng.tensor(phases, torch.float64), eng.tensor(amp, torch.float64)
    got = {}
    for mode in ("linear", "direct"):
        scr = screen(mode)
        slab = eng.empty((ncol, 4, n, n), torch.float32)
        scr._fill_columns(cols, slab, 0.1, None, sigma, True, amp=amps)
        got[mode] = slab.cpu().numpy()
        err = np.abs(got[mode] - want).max()
        assert err < 2 * TOL_SMOOTH, (mode, err)
    assert screen("auto")._use_linear(sigma, cols) == (n * n >= 128 * 128)
    # a non-finite amplitude sends the automatic choice to the direct form
    bad = screen("auto")
    bad._tmpl_dev = eng.tensor(np.ones((128, 128), np.int32), torch.int32)
    assert bad._use_linear(sigma, cols)
    amp_bad = amp.copy()
    amp_bad[0, 0, 0] = np.inf
    bad = screen("auto")
    bad._tmpl_dev = eng.tensor(np.ones((128, 128), np.int32), torch.int32)
    bad._dev_amp = eng.tensor(amp_bad, torch.float64)
    assert not bad._use_linear(sigma, cols)
    return float(np.abs(got["linear"] - got["direct"]).max())


def gain_solutions(golden_dir, repo_solutions):
    """{soltab name: Solutions} of the gain fixture: the repo's phase000 plus the synthetic
    amplitude000 stored in tests/golden/gain_c1.npz (oracle/make_golden_synth.py)."""
    import os

    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions

    gold = np.load(os.path.join(golden_dir, "gain_c1.npz"))
    ph = repo_solutions
    amp = Solutions("amplitude", gold["amp_val"], gold["amp_weight"], gold["amp_time"], gold["amp_freq"],
                    list(ph.ant), list(ph.dir), ["XX", "YY"], dict(ph.source_dict), dict(ph.station_dict))
    return {"phase000": ph, "amplitude000": amp}, gold


def check_voronoi_gain_golden(eng, golden_dir, repo_solutions, tmp_path, full=True):
    """VoronoiScreen on gain solutions through process() + write(): against the cube the
    reference's own code produced (smoothed and unsmoothed)."""
    from ska_sdp_screen_fitting_b200 import fits
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    sols, gold = gain_solutions(golden_dir, repo_solutions)
    import os

    for smooth_pix, key, tol in ((0.0, "cube_tessellated_nosmooth", 1e-30), (0.5, "cube_tessellated", TOL_SMOOTH)):
        scr = VoronoiScreen("gain", None, os.path.join(golden_dir, "skymodel.txt"), 126.23, 64.50, 3.33, 3.33,
                            phase_soltab_name="phase000", amplitude_soltab_name="amplitude000")
        scr._engine = eng
        scr.solutions = sols
        scr.process()
        if not full:
            # emulator-sized: two (freq, station) pairs through make_matrix, unsmoothed, bit-exact
            for fi, ai in ((0, 1), (2, 3)):
                one = scr.make_matrix(0, 20, int(gold["sub_f"][fi]), int(gold["sub_a"][ai]), 0.2, None, 0)
                assert np.array_equal(one.astype(np.float32), gold["cube_tessellated_nosmooth"][:, fi, ai])
            return
        scr.write(str(tmp_path), 0.2, smooth_pix=smooth_pix)
        _, data = fits.read(os.path.join(str(tmp_path), "gain_0.fits"))
        sub = np.asarray(data)[:, list(gold["sub_f"])][:, :, list(gold["sub_a"])]
        assert sub.shape == gold[key].shape
        assert np.abs(sub - gold[key]).max() <= tol, (key, np.abs(sub - gold[key]).max())
        one = scr.make_matrix(0, 20, int(gold["sub_f"][1]), int(gold["sub_a"][2]), 0.2, None, 0)
        if smooth_pix == 0.0:
            assert np.array_equal(one.astype(np.float32), gold[key][:, 1, 2])


def check_time_gaps_golden(eng, golden_dir, repo_solutions, tmp_path):
    """Screen.write on a time axis with two gaps: one FITS file per gap-free chunk, file list,
    per-file time cards and cube contents against what the reference wrote
    (tests/golden/gaps_c1.npz, oracle/make_golden_synth.py::golden_gaps)."""
    import os

    from ska_sdp_screen_fitting_b200 import fits
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    gold = np.load(os.path.join(golden_dir, "gaps_c1.npz"))
    ph = repo_solutions
    gapped = Solutions("phase", ph.val, ph.weight, gold["times"], ph.freq, list(ph.ant), list(ph.dir), [],
                       dict(ph.source_dict), dict(ph.station_dict))
    scr = VoronoiScreen("gap", None, os.path.join(golden_dir, "skymodel.txt"), 126.23, 64.50, 3.33, 3.33,
                        phase_soltab_name="phase001")
    scr._engine = eng
    scr.solutions = {"phase001": gapped}
    scr.process()
    outfiles = scr.write(str(tmp_path), 0.2, smooth_pix=0.0)
    assert [os.path.basename(f) for f in outfiles] == list(gold["files"])
    with open(os.path.join(str(tmp_path), "gap.txt"), encoding="utf8") as fh:
        assert [os.path.basename(ln.strip()) for ln in fh if ln.strip()] == list(gold["files"])
    for k, name in enumerate(outfiles):
        header, data = fits.read(name)
        assert header["NAXIS6"] == int(gold[f"naxis6_{k}"])
        assert header["CRVAL6"] == float(gold[f"crval6_{k}"]) and header["CDELT6"] == float(gold[f"cdelt6_{k}"])
        sub = np.asarray(data)[:, list(gold["sub_f"])][:, :, list(gold["sub_a"])]
        assert np.array_equal(sub, gold[f"cube_{k}"]), name


def check_kl_gain_golden(eng, golden_dir, repo_solutions, full=True):
    """KLScreen on gain solutions against what the reference's own code produced: the
    amplitude screen soltab (values, residuals, orders, weights) and the a-term cube."""
    import os

    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

    sols, gold = gain_solutions(golden_dir, repo_solutions)
    scr = KLScreen("gain", None, os.path.join(golden_dir, "skymodel.txt"), 126.23, 64.50, 3.33, 3.33,
                   phase_soltab_name="phase000", amplitude_soltab_name="amplitude000")
    scr._engine = eng
    scr.solutions = sols
    sub_a = [int(a) for a in gold["sub_a"]]
    if not full:
        # emulator-sized: the amplitude fit of the four fixture stations (stations are independent)
        scr.setup(sols["phase000"])
        amp = sols["amplitude000"]
        white, resid, order, w_out = scr.fit_amplitude_device(
            eng.tensor(amp.val[:, :, sub_a], torch.float64), eng.tensor(amp.weight[:, :, sub_a], torch.float32))
        assert np.abs(white.cpu().numpy() - gold["amp_screen"][:, :, sub_a]).max() < TOL_FIT
        assert np.abs(resid.cpu().numpy() - gold["amp_resid"][:, :, sub_a]).max() < TOL_FIT
        assert np.array_equal(order.cpu().numpy(), gold["amp_order"][:, :, sub_a])
        assert np.array_equal(w_out.cpu().numpy(), gold["amp_screen_weights"][:, :, sub_a])
        return
    scr.fit()
    assert np.abs(scr.vals_amp - gold["amp_screen"]).max() < TOL_FIT
    assert np.abs(scr.amp_residuals - gold["amp_resid"]).max() < TOL_FIT
    assert np.array_equal(scr.amp_screen_order, gold["amp_order"])
    assert np.array_equal(scr.amp_weights_out, gold["amp_screen_weights"])
    scr.interpolate()
    assert scr.vals_amp.shape[:2] == scr.vals_ph.shape[:2]
    for fi, f in enumerate(gold["sub_f"]):
        for ai, a in enumerate(sub_a):
            got = scr.make_matrix(0, 20, int(f), a, 0.2, None, 0)
            want = gold["cube_kl"][:, fi, ai].astype(np.float64)
            assert (np.abs(got - want) <= TOL_KL * np.maximum(1.0, np.abs(want))).all(), (fi, ai)


def check_smooth(eng, shape=(5, 19, 37), sigmas=(0.5, 2.0, 5.0), seed=2):
    rng = np.random.default_rng(seed)
    img = rng.normal(size=shape).astype(np.float32)
    worst = 0.0
    for sigma in sigmas:
        got = eng.gauss_smooth(eng.tensor(img, torch.float32), sigma).cpu().numpy()
        want = screens.smooth_block(img, sigma)
        worst = max(worst, float(np.abs(got - want).max()))
        assert np.abs(got - want).max() < TOL_SMOOTH * max(1.0, np.abs(img).max())
    return worst


def check_smoothed_voronoi(eng, n=24, n_d=6, ncol=5, sigma=1.5, seed=4, variant=0):
    rng = np.random.default_rng(seed)
    px, py = rng.uniform(0, n, n_d), rng.uniform(0, n, n_d)
    want_t = screens.nearest_patch_template(px, py, n)
    tmpl = eng.tensor(want_t, torch.int32)
    phases = rng.normal(0, 1.0, (1, 1, ncol, n_d))
    want = screens.voronoi_cube(phases, want_t, sigma)[0, 0]
    lut = eng.voronoi_lut(eng.tensor(phases.reshape(ncol, n_d), torch.float64))
    slab = eng.empty((ncol, 4, n, n), torch.float32)
    eng.voronoi_fill(tmpl, lut, slab, False)
    eng.gauss_smooth(slab, sigma)
    eng.nan_fix(slab, ncol, n * n)
    assert np.abs(slab.cpu().numpy() - want).max() < TOL_SMOOTH
    # a NaN phase spreads over the filter support and is then fixed (screen.py:353-378): the
    # fused form (ssf_gauss_smooth_slab) against smooth + separate NaN fix, bit for bit, and
    # against the oracle on the pixels the NaN does not reach
    ph_nan = phases.reshape(ncol, n_d).copy()
    ph_nan[0, 0] = np.nan
    lut_nan = eng.voronoi_lut(eng.tensor(ph_nan, torch.float64))
    a, b = eng.empty((ncol, 4, n, n), torch.float32), eng.empty((ncol, 4, n, n), torch.float32)
    eng.voronoi_fill(tmpl, lut_nan, a, False)
    eng.voronoi_fill(tmpl, lut_nan, b, False)
    eng.gauss_smooth(a, sigma)
    eng.nan_fix(a, ncol, n * n)
    eng.gauss_smooth_slab(b, sigma, fix_nan=True)
    assert np.array_equal(a.cpu().numpy(), b.cpu().numpy())
    want_nan = screens.voronoi_cube(ph_nan.reshape(1, 1, ncol, n_d), want_t, sigma)[0, 0]
    assert np.abs(b.cpu().numpy() - want_nan).max() < TOL_SMOOTH
    ind = eng.gauss_smooth(eng.voronoi_indicators(tmpl, n_d), sigma)
    slab2 = eng.empty((ncol, 4, n, n), torch.float32)
    eng.voronoi_contract(ind, lut, slab2, variant)
    assert np.abs(slab2.cpu().numpy() - want).max() < 2 * TOL_SMOOTH
    expect_formats(slab2.cpu().numpy(), lambda flags, planes: eng.voronoi_contract(
        ind, lut, eng.empty((ncol, planes, n, n), torch.float32), variant, flags).cpu().numpy())


def check_bswap_nanfix(eng, seed=1):
    rng = np.random.default_rng(seed)
    for count in (1, 7, 4096 + 3):
        a = rng.normal(size=count).astype(np.float32)
        t = eng.tensor(a, torch.float32)
        eng.bswap32(t)
        assert np.array_equal(t.cpu().numpy().view(np.uint32), a.view(np.uint32).byteswap())
    slab = rng.normal(size=(3, 4, 5, 5)).astype(np.float32)
    slab[rng.random(slab.shape) < 0.2] = np.nan
    t = eng.tensor(slab, torch.float32)
    eng.nan_fix(t, 3, 25)
    assert np.array_equal(t.cpu().numpy(), screens.nan_fix(slab.copy()))


def check_raw_contraction(eng, npix=1024, n_d=45, ncol=70, seed=11, variant=1):
    """The contraction alone (no sin/cos): random O(1) operands through ssf_voronoi_contract
    against a float64 matmul.  Bounds the split-TF32 error of the tensor-core variant by the
    same figure an FP32 FMA chain achieves (a few ulp of sum |a||b|)."""
    rng = np.random.default_rng(seed)
    s_mat = rng.normal(0, 1.0, (n_d, npix)).astype(np.float32)
    lut = rng.normal(0, 1.0, (ncol, n_d, 2)).astype(np.float32)
    lut = np.concatenate([lut, lut[-1:]])                    # extra column with one NaN coefficient:
    lut[-1, 0, 0] = np.nan                                   # only that column's Re becomes (1, .)
    ncol += 1
    slab = eng.empty((ncol, 4, npix), torch.float32)
    eng.voronoi_contract(eng.tensor(s_mat, torch.float32), eng.tensor(lut, torch.float32), slab, variant)
    got = slab.cpu().numpy()
    want_re = lut[:, :, 0].astype(np.float64) @ s_mat.astype(np.float64)
    want_im = lut[:, :, 1].astype(np.float64) @ s_mat.astype(np.float64)
    scale = np.abs(lut[:-1, :, 0]).astype(np.float64) @ np.abs(s_mat).astype(np.float64)
    err_re = np.abs(got[:-1, 0] - want_re[:-1]) / scale
    scale_im = np.abs(lut[:, :, 1]).astype(np.float64) @ np.abs(s_mat).astype(np.float64)
    err_im = np.abs(got[:, 1] - want_im) / scale_im
    assert err_re.max() < 4e-7, err_re.max()
    assert err_im.max() < 4e-7, err_im.max()
    assert np.all(got[-1, 0] == 1.0)                          # NaN column -> NaN fix (1, .)
    assert np.array_equal(got[:, 0], got[:, 2]) and np.array_equal(got[:, 1], got[:, 3])
    return float(err_re.max())


def check_streamed_write(eng, tmp_path, n_pix=16, shape="tiny", n_times=None, smooth=1.5, cellsize=None):
    """The API path: ``process()`` + ``write()`` streams (KL: the fit of the next block of time
    slots runs beside the evaluation; all screens: fused big-endian / plane-pair stores, D2H on a
    copy stream, parallel gathered file writes).  The cube must equal, byte for byte, the one
    written after an eager fit through native 4-plane slabs and the separate swap pass -- and the
    fit results kept from the streamed run must equal the eager ones."""
    import synthetic_cases
    from ska_sdp_screen_fitting_b200 import _lib, fits
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    case = synthetic_cases.make_case(shape, seed=11, n_pix=n_pix, n_times=n_times, cellsize_deg=cellsize)
    case.solutions.time[-2:] += 300.0                       # a gap: two output files
    n_f, n_a = len(case.solutions.freq), len(case.solutions.ant)

    def build(cls, name):
        scr = cls(name, None, None, case.rad, case.dec, case.width_deg, case.width_deg)
        scr._engine = eng
        scr.solutions = case.solutions
        if cls is VoronoiScreen:
            scr.patch_radec = case.patch_radec
            scr.template_mode = "nearest"
        scr.process()
        return scr

    def same_files(a, b):
        for fa, fb in zip(a, b):
            with open(fa, "rb") as ha, open(fb, "rb") as hb:
                assert ha.read() == hb.read(), (fa, fb)

    out = str(tmp_path)
    os.makedirs(out, exist_ok=True)
    # --- KL: streamed (blocks of 2 time slots, slabs of 5 columns) against eager + native slabs
    kl = build(KLScreen, "kl_streamed")
    kl.block_columns = 2 * n_f * n_a
    kl.slab_bytes = 5 * 8 * n_pix * n_pix
    assert kl._fitted is None
    files = kl.write(out, case.cellsize_deg)
    assert kl.timings["planes_copied"] == 2 and kl.timings["out_flags"] == 3
    assert kl.timings["d2h_bytes"] * 2 == kl.timings["file_bytes"] == kl.timings["voxels_written"] * 16
    eager = build(KLScreen, "kl_eager")
    eager.fit_now()
    eager._out_format = lambda smooth_pix: (0, 4)
    files_e = eager.write(out, case.cellsize_deg)
    assert eager.timings["planes_copied"] == 4 and len(files) == 2
    same_files(files, files_e)
    for name in ("vals_ph", "residuals", "screen_order", "weights_out"):
        assert np.array_equal(getattr(kl, name), getattr(eager, name)), name
    header, data = fits.read(files[0])
    one = eager.make_matrix(0, 2, 1, 2, case.cellsize_deg, None, 0)
    assert np.array_equal(np.asarray(data)[0:2, 1, 2], one.astype(np.float32))
    # --- Voronoi: unsmoothed fill and both smoothing forms against native slabs
    for tag, sm, mode in (("fill", 0.0, "auto"), ("linear", smooth, "linear"), ("direct", smooth, "direct")):
        vo = build(VoronoiScreen, f"vo_{tag}")
        vo.smooth_mode = mode
        vo.slab_bytes = 5 * 8 * n_pix * n_pix
        files = vo.write(out, case.cellsize_deg, smooth_pix=sm)
        assert vo.timings["planes_copied"] == (4 if tag == "direct" else 2)
        ref = build(VoronoiScreen, f"vo_{tag}_native")
        ref.smooth_mode = mode
        ref._out_format = lambda smooth_pix: (0, 4)
        same_files(files, ref.write(out, case.cellsize_deg, smooth_pix=sm))
    # --- "null" sink: the same pipeline up to the pinned host buffer, no file
    kl2 = build(KLScreen, "kl_null")
    kl2.sink = "null"
    kl2.write(out, case.cellsize_deg)
    assert kl2.timings["file_bytes"] == 0 and kl2.timings["voxels_written"] == kl.timings["voxels_written"]
    assert not os.path.exists(os.path.join(out, "kl_null_0.fits"))


def phase_golden_solutions(golden_dir, tag):
    """(Solutions, golden arrays) of tests/golden/phase45.npz / phase100.npz: the reference's own
    make_aterm_image(screen_type="kl") on a seeded 45- / 100-direction case with flags, outliers
    the 5-sigma circular clip fires on and orders that move (oracle/make_golden_synth.py::golden_phase)."""
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions

    g = np.load(os.path.join(golden_dir, f"{tag}.npz"))
    sol = Solutions.from_arrays(g["val"], g["weight"], g["time"], g["freq"], [str(a) for a in g["ant"]],
                                [str(d) for d in g["dir"]], g["source_rad"], g["station_xyz"])
    assert (g["screen_weights"] != g["weight"]).any(), "the fixture must hold clipped solutions"
    return sol, g


def check_phase_golden(eng, golden_dir, tag, cube=True):
    """The CUDA path at real direction counts against the reference itself: fit (white screens,
    residuals, weights after the 5-sigma clip, final orders) and, with ``cube``, the evaluated
    a-term planes of every (time, freq, station) at the fixture's 8 x 8 pixel sample of the
    512 x 512 grid."""
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

    sol, g = phase_golden_solutions(golden_dir, tag)
    scr = KLScreen("g", None, None, float(g["rad"]), float(g["dec"]), float(g["width_deg"]), float(g["width_deg"]))
    scr._engine = eng
    scr.solutions = sol
    scr.process()
    assert np.abs(scr.piercepoints - g["piercepoints"]).max() < 1e-9
    assert np.array_equal(scr.weights_out, g["screen_weights"])
    assert np.array_equal(scr.screen_order, g["order"])
    assert np.abs(scr.vals_ph - g["screen"]).max() < TOL_FIT
    assert np.abs(scr.residuals - g["resid"]).max() < TOL_FIT
    if not cube:
        return
    ix = g["pix"]
    n_t, n_f, n_a = g["order"].shape
    for f in range(n_f):
        for a in range(n_a):
            got = scr.make_matrix(0, n_t, f, a, float(g["cellsize_deg"]), None, 0)
            assert got.shape[-1] == 512
            want = g["cube_sample"][:, f, a].astype(np.float64)
            assert np.abs(got[..., ix[:, None], ix[None, :]] - want).max() < TOL_KL, (f, a)


def check_soltab_write_back(eng, golden_dir, tmp_path, sel_f=(0, 11), sel_a=(0, 1, 30, 61), full=False):
    """``write_soltabs``: after process() + write() the h5parm holds ``phase_screen000`` /
    ``phase_screen000resid`` as the reference leaves them (stationscreen.py:1087-1159): values,
    float16 weights (flags / orders), AXES, attributes, pierce points -- read back with the h5
    reader and compared with the in-memory results and, where the station orders agree with the
    full array's, with what the reference wrote (tests/golden/kl_c2.npz).  The input soltab is a
    sub-cube appended to a copy of the repo file with the package's own appender, unless ``full``."""
    import shutil

    from ska_sdp_screen_fitting_b200.h5parm_io import read_solutions
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen, station_orders
    from ska_sdp_screen_fitting_b200.utils import h5min, h5write

    gold = np.load(os.path.join(golden_dir, "kl_c2.npz"))
    path = os.path.join(str(tmp_path), "solutions_wb.h5")
    os.makedirs(str(tmp_path), exist_ok=True)
    shutil.copy(os.path.join(golden_dir, "solutions.h5"), path)
    whole = read_solutions(path, "sol000", "phase000")
    soltab = "phase000"
    sel_f, sel_a = (list(range(len(whole.freq))), list(range(len(whole.ant)))) if full else (list(sel_f), list(sel_a))
    if not full:
        soltab = "phase001"
        h5write.append_soltab(path, "sol000", soltab, "phase", ["time", "freq", "ant", "dir"],
                              [whole.time, whole.freq[sel_f], np.array([whole.ant[a].encode() for a in sel_a]),
                               np.array([d.encode() for d in whole.dir])],
                              whole.val[:, sel_f][:, :, sel_a], whole.weight[:, sel_f][:, :, sel_a])
    scr = KLScreen("wb", path, os.path.join(golden_dir, "skymodel.txt"), 126.23, 64.50, 3.33, 3.33,
                   phase_soltab_name=soltab)
    scr._engine = eng
    scr.write_soltabs = True
    scr.process()
    scr.write(str(tmp_path), 0.2)
    with h5min.File(path) as handle:
        assert np.array_equal(handle["sol000/phase000/val"].read().reshape(whole.val.shape), whole.val)   # untouched
        tab, res = handle["sol000/phase_screen000"], handle["sol000/phase_screen000resid"]
        assert tab.attrs["TITLE"] == "phasescreen" and res.attrs["TITLE"] == "phasescreenresid"
        assert tab["val"].attrs["AXES"] == "time,freq,ant,dir" == res["weight"].attrs["AXES"]
        assert tab["weight"].dtype == np.float16
        for key, want in (("beta", 5.0 / 3.0), ("r_0", 100.0), ("height", 0.0), ("midra", float(gold["midra"])),
                          ("middec", float(gold["middec"]))):
            assert abs(float(tab.attrs[key]) - want) < 1e-9, key
        assert np.abs(tab["piercepoint"].read() - gold["piercepoints"]).max() < 1e-9
        white, w_out = tab["val"].read(), tab["weight"].read()
        resid, order = res["val"].read(), res["weight"].read()
        assert [a.decode() for a in tab["ant"].read()] == [whole.ant[a] for a in sel_a]
    assert np.array_equal(white, scr.vals_ph) and np.array_equal(resid, scr.residuals)
    assert np.array_equal(w_out, scr.weights_out.astype(np.float16))
    assert np.array_equal(order, np.broadcast_to(scr.screen_order[..., None], order.shape).astype(np.float16))
    pos = np.asarray(whole.station_positions)
    orders_full = station_orders(pos, 0, 6)
    same = [i for i, a in enumerate(sel_a) if station_orders(pos[sel_a], 0, 6)[i] == orders_full[a]]
    assert len(same) >= 2
    for i in same:
        f_ix, a = np.array(sel_f), sel_a[i]
        assert np.abs(white[:, :, i] - gold["screen"][:, f_ix, a]).max() < TOL_FIT
        assert np.abs(resid[:, :, i] - gold["resid"][:, f_ix, a]).max() < TOL_FIT
        assert np.array_equal(order[:, :, i, 0], gold["order"][:, f_ix, a])
        assert np.array_equal(w_out[:, :, i], gold["screen_weights"][:, f_ix, a])
    # the table just written is a valid INPUT: the reader serves it back as a soltab
    back = read_solutions(path, "sol000", "phase_screen000")
    assert back.val.shape == white.shape and back.ant == [whole.ant[a] for a in sel_a]


def check_tec_run_golden(eng, golden_dir, tmp_path):
    """A soltab of type "tec" through the public classes: KLScreen.process() fits it with the tec
    branches of the operators (against the reference's own stationscreen.run on that soltab) and
    write() makes the 5-axis dTEC cube [t, f, a, y, x] of processing_utils.py:187-190 (against the
    screen images of the reference's calculate_kl_screen) -- tests/golden/tec_run.npz."""
    from ska_sdp_screen_fitting_b200 import fits
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

    g = np.load(os.path.join(golden_dir, "tec_run.npz"))
    n_t, n_f, n_a, n_d = g["val"].shape
    sol = Solutions.from_arrays(g["val"], g["weight"], g["time"], g["freq"], [f"ST{i:03d}" for i in range(n_a)],
                                [f"[Patch_{i}]" for i in range(n_d)], g["dirs"], g["station_xyz"], soltype="tec")
    n_pix, cell = int(g["n_pix"]), float(g["cellsize_deg"])
    scr = KLScreen("tec", None, None, float(g["rad"]), float(g["dec"]), (n_pix - 0.5) * cell, (n_pix - 0.5) * cell)
    scr._engine = eng
    scr.solutions = sol
    scr.process()
    assert scr.aterm_type == "tec" and scr.ref_ind == int(g["ref_ind"])
    assert np.array_equal(scr.weights_out, g["screen_weights"])
    assert (g["screen_weights"] != g["weight"]).any(), "the fixture must hold clipped solutions"
    assert np.array_equal(scr.screen_order, g["order"])
    assert np.abs(scr.vals_ph - g["screen"]).max() < TOL_FIT
    assert np.abs(scr.residuals - g["resid"]).max() < TOL_FIT * 20
    os.makedirs(str(tmp_path), exist_ok=True)
    files = scr.write(str(tmp_path), cell)
    header, data = fits.read(files[0])
    assert header["NAXIS"] == 5 and data.shape == (n_t, n_f, n_a, n_pix, n_pix)
    assert [header[f"CTYPE{i}"] for i in range(1, 6)] == ["RA---SIN", "DEC--SIN", "ANTENNA", "FREQ", "TIME"]
    scale = max(1.0, float(np.abs(g["images"]).max()))
    assert np.abs(np.asarray(data, np.float64) - g["images"]).max() < TOL_KL * scale
    one = scr.make_tec_matrix(0, n_t, 1, 2, cell)
    assert np.array_equal(one.astype(np.float32), np.asarray(data)[:, 1, 2])

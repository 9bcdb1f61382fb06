"""Parity of the CUDA path (libssf_b200.so on a real GPU, through the C ABI) with the oracle."""
import os
import shutil

import numpy as np
import pytest
import torch

import parity_checks as pc

pytestmark = pytest.mark.gpu


def test_decompose(gpu_engine):
    pc.check_decompose(gpu_engine)
    pc.check_decompose(gpu_engine, n_d=45, seed=6)
    pc.check_decompose(gpu_engine, n_d=100, seed=7)


def test_pattern_cache_warm_started_decompositions(gpu_engine):
    pc.check_pattern_cache_decompose(gpu_engine)
    pc.check_pattern_cache_decompose(gpu_engine, n_d=45, seed=8, max_flagged=4, n_sys=12)
    pc.check_pattern_cache_decompose(gpu_engine, n_d=100, seed=9, max_flagged=5, n_sys=12)
    pc.check_pattern_cache_decompose(gpu_engine, n_d=99, seed=10, max_flagged=3, n_sys=8)
    pc.check_pattern_cache_decompose(gpu_engine, n_d=112, seed=11, max_flagged=6, n_sys=8)   # SSF_MAX_DIR
    pc.check_pattern_cache_decompose(gpu_engine, n_d=65, seed=12, max_flagged=2, n_sys=6)    # first 512-thread size
    pc.check_pattern_cache_decompose(gpu_engine, n_d=64, seed=13, max_flagged=2, n_sys=6)    # last 256-thread size


def test_pattern_cache_under_pressure(gpu_engine):
    pc.check_pattern_cache_pressure(gpu_engine)


def test_fit_tec_type_against_reference_operator_golden(gpu_engine, golden_dir):
    pc.check_fit_tec_golden(gpu_engine, golden_dir)


def test_tec_soltab_fit_and_dtec_cube_against_reference_golden(gpu_engine, golden_dir, tmp_path):
    pc.check_tec_run_golden(gpu_engine, golden_dir, tmp_path)


def test_time_gaps_give_one_file_per_chunk_like_the_reference(gpu_engine, golden_dir, repo_solutions, tmp_path):
    pc.check_time_gaps_golden(gpu_engine, golden_dir, repo_solutions, tmp_path)


def test_streamed_write_equals_eager_write(gpu_engine, tmp_path):
    """The API path on the real device: side-stream fits, copy-stream D2H, tcgen05 evaluation with
    fused big-endian plane-pair stores -- byte-identical files to the eager / native path."""
    pc.check_streamed_write(gpu_engine, tmp_path)
    pc.check_streamed_write(gpu_engine, tmp_path / "b", n_pix=128, shape="small", n_times=9, cellsize=0.025,
                            smooth=2.0)


def test_soltab_write_back(gpu_engine, golden_dir, tmp_path):
    """The whole repo data set: the soltabs appended to the h5parm against the ones the reference wrote."""
    pc.check_soltab_write_back(gpu_engine, golden_dir, tmp_path, full=True)


def test_fit_with_weights_below_the_pinv_cutoff(gpu_engine):
    pc.check_fit_small_weights(gpu_engine)


def test_fit_with_flags_and_outliers(gpu_engine):
    pc.check_fit(gpu_engine)
    pc.check_fit(gpu_engine, seed=17, n_t=6, n_f=3, n_a=8, n_d=45)


def test_fit_repo_data_against_reference_golden(gpu_engine, repo_solutions, golden_dir):
    """BASELINE.json configs[1]: the whole repo data set, all 14 880 systems."""
    gold = np.load(os.path.join(golden_dir, "kl_c2.npz"))
    sol = repo_solutions
    scr = pc.run_engine_fit(gpu_engine, sol.val, sol.weight, np.asarray(sol.source_positions),
                            np.asarray(sol.station_positions))
    assert np.abs(scr.piercepoints - gold["piercepoints"]).max() < 1e-9
    assert np.abs(scr.vals_ph - gold["screen"]).max() < pc.TOL_FIT
    assert np.abs(scr.residuals - gold["resid"]).max() < pc.TOL_FIT
    assert np.array_equal(scr.screen_order, gold["order"])
    assert np.array_equal(scr.weights_out, gold["screen_weights"])


@pytest.mark.parametrize("tag", ["phase45", "phase100"])
def test_fit_and_screens_at_real_direction_counts_against_the_reference(gpu_engine, golden_dir, tag):
    """BASELINE.json configs[2] / configs[4] direction counts: triggered 5-sigma clip,
    re-decomposition of the flagged subsets, order adjustment and the 512 x 512 evaluation, CUDA
    against outputs of the reference's own code."""
    pc.check_phase_golden(gpu_engine, golden_dir, tag)


def test_kl_eval(gpu_engine):
    pc.check_kl_eval(gpu_engine, n=12, ncol=9)
    pc.check_kl_eval(gpu_engine, n=11, n_d=5, ncol=3, seed=8)       # ragged: npix % 4 != 0
    pc.check_kl_eval(gpu_engine, n=64, n_d=45, ncol=70, seed=9)
    pc.check_kl_eval(gpu_engine, n=48, n_d=100, ncol=33, seed=10)


def test_kl_eval_tcgen05(gpu_engine):
    """Variant 1 (tcgen05, split TF32) against the float64 oracle at the same 1e-5 tolerance:
    K padding (9 -> 16, 45 -> 48, 100 -> 104, 112), ragged column tiles, 1 and 2 CTAs per SM."""
    pc.check_kl_eval(gpu_engine, n=32, n_d=9, ncol=11, variant=1)
    pc.check_kl_eval(gpu_engine, n=64, n_d=45, ncol=70, seed=9, variant=1)
    pc.check_kl_eval(gpu_engine, n=48, n_d=100, ncol=133, seed=10, variant=1)
    pc.check_kl_eval(gpu_engine, n=32, n_d=112, ncol=64, seed=14, variant=1)
    pc.check_kl_eval(gpu_engine, n=128, n_d=45, ncol=700, seed=15, variant=1, cell=0.03)
    # many column tiles per CTA (both TMEM / shared-memory ring stages reused many times)
    pc.check_kl_eval(gpu_engine, n=128, n_d=45, ncol=3840, seed=16, variant=1, cell=0.03)
    # more than 56 directions -> persistent kernel: several items per CTA (basis tile prefetch and
    # re-split, all ring phases wrap), ragged last column tile, column chunks
    pc.check_kl_eval(gpu_engine, n=160, n_d=100, ncol=77, seed=17, variant=1, cell=0.02)
    pc.check_kl_eval(gpu_engine, n=256, n_d=60, ncol=300, seed=18, variant=1, cell=0.015)
    pc.check_kl_eval(gpu_engine, n=64, n_d=112, ncol=1500, seed=19, variant=1, cell=0.05)


def test_contraction_tcgen05_error_bound(gpu_engine):
    for n_d, ncol in ((45, 70), (8, 32), (100, 33), (3, 1)):
        pc.check_raw_contraction(gpu_engine, n_d=n_d, ncol=ncol, variant=1)
    pc.check_raw_contraction(gpu_engine, npix=128 * 37, n_d=45, ncol=500, variant=1)
    # persistent TMEM-operand kernel, Voronoi rows (Re/Im), several items per CTA, ragged tile
    pc.check_raw_contraction(gpu_engine, npix=128 * 200, n_d=64, ncol=100, seed=12, variant=1)
    pc.check_raw_contraction(gpu_engine, n_d=45, ncol=70, variant=0)


def test_tcgen05_rejects_unsupported_shapes(gpu_engine):
    from ska_sdp_screen_fitting_b200._lib import SsfError
    with pytest.raises(SsfError):
        pc.check_kl_eval(gpu_engine, n=11, n_d=5, ncol=3, seed=8, variant=1)   # npix % 128 != 0


def test_voronoi_assign_and_fill(gpu_engine):
    pc.check_voronoi(gpu_engine)
    pc.check_voronoi(gpu_engine, n=9, n_d=3, ncol=2, seed=12)
    pc.check_voronoi(gpu_engine, n=128, n_d=45, ncol=40, seed=13)


def test_kl_gain(gpu_engine, golden_dir, repo_solutions):
    pc.check_fit_amplitude(gpu_engine)
    pc.check_fit_amplitude(gpu_engine, seed=29, n_t=9, n_f=3, n_a=6, n_d=45)
    pc.check_kl_gain_eval(gpu_engine)
    pc.check_kl_gain_eval(gpu_engine, n=11, n_d=5, ncol=3, seed=8)
    pc.check_kl_gain_eval(gpu_engine, n=128, n_d=45, ncol=33, seed=9, cell=0.02)
    # tcgen05 variant: ragged column tiles (16 per tile), one- and two-CTA-per-SM shapes
    pc.check_kl_gain_eval(gpu_engine, n=128, n_d=45, ncol=33, seed=9, cell=0.02, variant=1)
    pc.check_kl_gain_eval(gpu_engine, n=64, n_d=9, ncol=7, seed=10, cell=0.05, variant=1)
    pc.check_kl_gain_eval(gpu_engine, n=64, n_d=100, ncol=50, seed=11, cell=0.05, variant=1)
    pc.check_kl_gain_eval(gpu_engine, n=160, n_d=100, ncol=21, seed=12, cell=0.02, variant=1)   # persistent, multi-item
    pc.check_kl_gain_golden(gpu_engine, golden_dir, repo_solutions)


def test_voronoi_gain(gpu_engine, golden_dir, repo_solutions, tmp_path):
    pc.check_voronoi_gain(gpu_engine)
    pc.check_voronoi_gain(gpu_engine, n=9, n_d=3, ncol=2, seed=12)
    pc.check_voronoi_gain(gpu_engine, n=256, n_d=45, ncol=40, seed=13)
    pc.check_voronoi_gain_golden(gpu_engine, golden_dir, repo_solutions, tmp_path)


def test_template_bit_exact_at_full_size(gpu_engine):
    """1024 x 1024, 45 patches: every index equals the float64 CPU definition."""
    from oracle.port import screens

    rng = np.random.default_rng(4)
    px, py = rng.uniform(100, 900, 45), rng.uniform(100, 900, 45)
    tmpl = gpu_engine.voronoi_assign(gpu_engine.tensor(px, torch.float64),
                                     gpu_engine.tensor(py, torch.float64), 1024, 1024)
    assert np.array_equal(tmpl.cpu().numpy(), screens.nearest_patch_template(px, py, 1024))


@pytest.mark.parametrize("n", [512, 1024])
def test_nearest_patch_kernel_against_the_reference_template(gpu_engine, n, capsys):
    """SURVEY 7.4-1 iii: the two definitions of "the Voronoi template" side by side.  The
    reference rasterises Voronoi polygons with PIL (integer vertex truncation, host pipeline
    restated in template.py = the default template_mode="reference"); ``assign_kernel`` is the
    exact float64 nearest-patch rule.  They may only differ along cell boundaries: the count is
    printed (and bounded by the total boundary length)."""
    import synthetic_cases
    from ska_sdp_screen_fitting_b200 import template, wcs

    case = synthetic_cases.make_case("lofar", seed=3, n_times=1, n_freqs=1, n_pix=n, cellsize_deg=5.12 / n)
    ra, dec = case.patch_radec[:, 0], case.patch_radec[:, 1]
    ref, _, _ = template.reference_template(ra, dec, case.rad, case.dec, case.width_deg, case.cellsize_deg)
    assert ref.shape == (n, n)
    px, py = wcs.patch_pixels(ra, dec, case.rad, case.dec, n, case.cellsize_deg)
    got = gpu_engine.voronoi_assign(gpu_engine.tensor(px, torch.float64), gpu_engine.tensor(py, torch.float64),
                                    n, n).cpu().numpy()
    differ = got != ref
    count = int(differ.sum())
    # every differing pixel touches a cell boundary of the exact template (8-neighbourhood)
    edge = np.zeros_like(differ)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            shifted = np.roll(np.roll(got, dy, 0), dx, 1)
            edge |= shifted != got
    with capsys.disabled():
        print(f"\n[template] {n} x {n}, {len(ra)} patches: assign_kernel differs from the reference-style template "
              f"in {count} of {n * n} pixels ({100.0 * count / (n * n):.3f} %), boundary pixels {int(edge.sum())}")
    assert not (differ & ~edge).any()
    assert count <= edge.sum()


def test_smoothing(gpu_engine):
    pc.check_smooth(gpu_engine)
    pc.check_smooth(gpu_engine, shape=(3, 130, 257), sigmas=(0.5, 2.0, 5.0, 10.0), seed=5)


def test_smoothed_voronoi_both_forms(gpu_engine):
    pc.check_smoothed_voronoi(gpu_engine)
    pc.check_smoothed_voronoi(gpu_engine, n=96, n_d=20, ncol=9, sigma=5.0, seed=6)
    pc.check_smoothed_voronoi(gpu_engine, n=96, n_d=20, ncol=9, sigma=5.0, seed=6, variant=1)
    pc.check_smoothed_voronoi(gpu_engine, n=64, n_d=45, ncol=100, sigma=2.0, seed=7, variant=1)


def test_smoothed_gain_voronoi_linear_form(gpu_engine):
    """Gain slabs through the tcgen05 plane-pair contraction (variant auto: 128 x 128 is the
    smallest image it takes) and through the SIMT contraction, against the float64 oracle."""
    pc.check_smoothed_voronoi_gain(gpu_engine)
    diff = pc.check_smoothed_voronoi_gain(gpu_engine, n=128, n_d=45, n_t=20, sigma=2.5, seed=29)
    assert diff < 2 * pc.TOL_SMOOTH


def test_bswap_and_nanfix(gpu_engine):
    pc.check_bswap_nanfix(gpu_engine)


def test_full_size_properties(gpu_engine):
    """BASELINE-size slab (512^2, 45 directions): size-independent properties instead of an
    oracle run -- planes 0 == 2 and 1 == 3, cos^2 + sin^2 == 1, linearity of the phase in the
    coefficients (a column with doubled white has doubled phase), no NaN."""
    eng = gpu_engine
    rng = np.random.default_rng(3)
    from oracle.port import klfit
    n, n_d, ncol = 512, 45, 32
    pp, mid_ra, mid_dec = klfit.piercepoints(pc.make_directions(rng, n_d, 2.0))
    from ska_sdp_screen_fitting_b200 import wcs
    x, y = wcs.kl_grid(126.23, 64.5, (n - 0.5) * 0.01, 0.01, mid_ra, mid_dec)
    ppt = eng.tensor(pp[:, :2], torch.float64)
    full = torch.tensor([pc.mask_of(range(n_d))], dtype=torch.int64, device=eng.device)
    c, p, _, _ = eng.kl_decompose(ppt, full)
    fit = rng.normal(0, 0.3, (ncol, n_d))
    fit[1] = 2.0 * fit[0]
    white = eng.tensor(fit, torch.float64) @ p[0].T
    bt = eng.kl_basis(eng.tensor(x, torch.float64), eng.tensor(y, torch.float64), ppt, p[0].contiguous())
    g = eng.kl_coeffs(white.contiguous(), c[0].contiguous())
    slab = eng.empty((ncol, 4, n, n), torch.float32)
    eng.kl_eval(bt, g, n_d, slab, -1)
    assert torch.equal(slab[:, 0], slab[:, 2]) and torch.equal(slab[:, 1], slab[:, 3])
    assert not torch.isnan(slab).any()
    assert float((slab[:, 0] ** 2 + slab[:, 1] ** 2 - 1).abs().max()) < 1e-5
    ph0 = torch.atan2(slab[0, 1].double(), slab[0, 0].double())
    ph1 = torch.atan2(slab[1, 1].double(), slab[1, 0].double())
    d = torch.remainder(ph1 - 2 * ph0 + np.pi, 2 * np.pi) - np.pi
    assert float(d.abs().max()) < 5e-5
    # sampled pixels against the float64 oracle formula K(pix, :) . white
    from oracle.port import screens
    sel = rng.integers(0, n, (64, 2))
    kmat = screens.kl_kernel_matrix(x[sel[:, 0]], y[sel[:, 1]], pp)       # [64, 64, d]
    want = np.einsum("iid,cd->ci", kmat, white.cpu().numpy())
    got_c = slab[:, 0].cpu().numpy()[:, sel[:, 1], sel[:, 0]]
    got_s = slab[:, 1].cpu().numpy()[:, sel[:, 1], sel[:, 0]]
    assert np.abs(got_c - np.cos(want)).max() < pc.TOL_KL
    assert np.abs(got_s - np.sin(want)).max() < pc.TOL_KL


def test_fill_kernels_beyond_one_grid_pass(gpu_engine):
    """The plane-pair and gain fills stride over a grid capped at 512 CTAs per SM: at 512^2 x 320
    columns their threads take more than one pass (the small parity cases never do), and the
    four-plane fill launches its largest one-pass grid.  Bit-exact against a torch gather of the
    same LUT, for the phase fill (4 planes, plane pairs big-endian) and the gain fill."""
    from ska_sdp_screen_fitting_b200 import _lib

    eng = gpu_engine
    gen = torch.Generator(device=eng.device)
    gen.manual_seed(17)
    n, n_d, ncol = 512, 45, 320
    npix = n * n
    tmpl = torch.randint(1, n_d + 1, (n, n), generator=gen, device=eng.device, dtype=torch.int32)
    idx = (tmpl.view(-1) - 1).long()
    lut = torch.randn((ncol, n_d, 2), generator=gen, device=eng.device, dtype=torch.float32)
    slab = eng.empty((ncol, 4, npix), torch.float32)
    eng.voronoi_fill(tmpl, lut, slab)
    want = lut[:, idx, :].permute(0, 2, 1)                      # [ncol, 2, npix]
    assert torch.equal(slab[:, 0], want[:, 0]) and torch.equal(slab[:, 1], want[:, 1])
    assert torch.equal(slab[:, 2], want[:, 0]) and torch.equal(slab[:, 3], want[:, 1])
    pair = eng.empty((ncol, 2, npix), torch.float32)
    eng.voronoi_fill(tmpl, lut, pair, out_flags=_lib.OUT_PLANES2 | _lib.OUT_BIG_ENDIAN)
    swapped = pair.view(torch.uint8).view(ncol, 2, npix, 4).flip(-1).contiguous().view(torch.float32).view(ncol, 2, npix)
    assert torch.equal(swapped, want)
    del pair, swapped
    lut4 = torch.randn((ncol, n_d, 4), generator=gen, device=eng.device, dtype=torch.float32)
    eng.voronoi_fill_gain(tmpl, lut4, slab)
    assert torch.equal(slab, lut4[:, idx, :].permute(0, 2, 1))


def test_host_entry_points(gpu_engine):
    import ctypes

    from oracle.port import klfit, screens, wcs as owcs
    from ska_sdp_screen_fitting_b200 import _lib

    rng = np.random.default_rng(21)
    n, n_d, ncol = 40, 9, 24
    tmpl = screens.nearest_patch_template(rng.uniform(0, n, n_d), rng.uniform(0, n, n_d), n)
    phases = rng.normal(size=(ncol, n_d))
    out = np.zeros((ncol, 4, n, n), np.float32)
    for sigma in (0.0, 1.0):
        _lib.call("ssf_voronoi_cube_host", tmpl.ctypes.data, phases.ctypes.data, n_d, n, n, ncol,
                  ctypes.c_double(sigma), out.ctypes.data)
        want = screens.voronoi_cube(phases.reshape(1, 1, ncol, n_d), tmpl, sigma)[0, 0]
        assert np.abs(out - want).max() < (pc.TOL_SMOOTH if sigma else 1e-30)
    pp, mid_ra, mid_dec = klfit.piercepoints(pc.make_directions(rng, n_d, 1.0))
    x, y = owcs.kl_grid_coords(126.23, 64.5, (n - 0.5) * 0.05, 0.05, mid_ra, mid_dec)
    _, p_o, _ = klfit.calculate_svd(pp)
    white = rng.normal(0, 0.5, (ncol, n_d)) @ p_o.T
    pp2 = np.ascontiguousarray(pp[:, :2])
    _lib.call("ssf_kl_cube_host", x.ctypes.data, y.ctypes.data, n, n, pp2.ctypes.data, n_d,
              white.ctypes.data, ncol, 100.0, 5.0 / 3.0, 1e-3, out.ctypes.data)
    assert np.abs(out - screens.kl_make_matrix(x, y, pp, white)).max() < pc.TOL_KL


def test_streamed_kl_slabs_match_the_whole_fit(gpu_engine):
    """KLScreen.stream_slabs (grouped fits on a side stream issued by a helper thread, ragged
    last group) gives the same cube slabs as one fit over all time slots followed by the
    evaluation -- the fit systems are independent across time (stationscreen.py:681-722)."""
    import synthetic_cases
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

    case = synthetic_cases.make_case("small", seed=11, n_times=10, flag_fraction=0.05)
    sol = case.solutions
    eng = gpu_engine
    kl = KLScreen("t", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    kl._engine = eng
    kl.setup(sol)
    val, wt = eng.tensor(sol.val, torch.float64), eng.tensor(sol.weight, torch.float32)
    n_t, n_f, n_a, n_d = sol.val.shape
    n, slots = case.n_pix, 2
    cols = slots * n_f * n_a
    white = kl.fit_device(val, wt)[0]
    want = []
    for k in range(n_t // slots):
        slab = eng.empty((cols, 4, n, n), torch.float32)
        kl._fill_columns(white[k * slots:(k + 1) * slots].reshape(-1, n_d), slab, case.cellsize_deg, None, 0.0, True)
        want.append(slab.cpu().numpy())
    for group in (1, 2, 4):
        kl.fit_group = group
        ring = [eng.empty((cols, 4, n, n), torch.float32) for _ in range(2)]
        seen = 0
        for k, slab in kl.stream_slabs(val, wt, slots, ring, case.cellsize_deg):
            assert k == seen
            got = slab.cpu().numpy()
            assert np.abs(got - want[k]).max() < 2e-6, (group, k)
            seen += 1
        assert seen == n_t // slots
        eng.check_fit_status()


@pytest.mark.parametrize("screen_type", ["tessellated", "kl"])
def test_make_aterm_image_repo_configs(gpu_engine, golden_dir, tmp_path, screen_type):
    """BASELINE.json configs[0] / configs[1] through the public API, against the cubes the
    reference's own code produced (tests/golden, oracle/make_golden.py) and the reference's
    own test assertions (tests/test_fit_screens.py:101-128, :190-215)."""
    from ska_sdp_screen_fitting_b200 import fits
    from ska_sdp_screen_fitting_b200.make_aterm_images import make_aterm_image

    for f in ("solutions.h5", "skymodel.txt"):
        shutil.copy(os.path.join(golden_dir, f), tmp_path)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        make_aterm_image("solutions.h5", soltabname="phase000", screen_type=screen_type,
                         outroot=screen_type, bounds_deg=[124.565, 66.165, 127.895, 62.835],
                         bounds_mid_deg=[126.23, 64.50], skymodel="skymodel.txt", solsetname="sol000",
                         padding_fraction=0, cellsize_deg=0.2, smooth_deg=0.1, ncpu=0)
        assert os.path.exists(f"{screen_type}_0.fits") and os.path.exists(f"{screen_type}.txt")
        header, data = fits.read(f"{screen_type}_0.fits")
    finally:
        os.chdir(cwd)
    assert data.shape == (20, 12, 62, 4, 17, 17)
    assert header["CTYPE1"] == "RA---SIN" and header["CTYPE3"] == "MATRIX" and header["CTYPE6"] == "TIME"
    if screen_type == "tessellated":
        assert os.path.exists(tmp_path / "tessellated_template.fits")
        gold = np.load(os.path.join(golden_dir, "voronoi_c1.npz"))
        want, tol = gold["cube_smooth"], pc.TOL_SMOOTH
    else:
        gold = np.load(os.path.join(golden_dir, "kl_c2.npz"))
        want, tol = gold["cube"], pc.TOL_KL
    sub = np.asarray(data)[:, list(gold["sub_f"])][:, :, list(gold["sub_a"])]
    assert np.abs(sub - want).max() < tol

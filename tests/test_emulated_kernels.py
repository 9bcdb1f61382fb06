"""Kernel LOGIC checked without a GPU: the product's .cu sources compiled by g++ against the
CUDA emulator (tests/emu) and compared with the oracle.  The same checks run on the real
library in test_gpu_parity.py."""
import numpy as np

import parity_checks as pc


def test_decompose(emu_engine):
    pc.check_decompose(emu_engine)


def test_pattern_cache_warm_started_decompositions(emu_engine):
    pc.check_pattern_cache_decompose(emu_engine)
    pc.check_pattern_cache_decompose(emu_engine, n_d=14, seed=8, max_flagged=5)
    # 3 of 64 / 65 kept: far beyond the warm-start limit (cold start), both thread-count variants
    pc.check_pattern_cache_decompose(emu_engine, n_d=64, seed=13, max_flagged=2, n_sys=6)
    pc.check_pattern_cache_decompose(emu_engine, n_d=65, seed=12, max_flagged=2, n_sys=6)


def test_pattern_cache_under_pressure(emu_engine):
    pc.check_pattern_cache_pressure(emu_engine)


def test_fit_with_weights_below_the_pinv_cutoff(emu_engine):
    pc.check_fit_small_weights(emu_engine)


def test_fit_with_flags_and_outliers(emu_engine):
    pc.check_fit(emu_engine)


def test_fit_repo_data_subset_against_reference_golden(emu_engine, repo_solutions, golden_dir):
    import os

    gold = np.load(os.path.join(golden_dir, "kl_c2.npz"))
    sol = repo_solutions
    sel_f, sel_a = [0, 11], [0, 1, 30, 61]
    val = sol.val[:, sel_f][:, :, sel_a]
    wt = sol.weight[:, sel_f][:, :, sel_a]
    pos = np.asarray(sol.station_positions)
    # station orders depend on the distance to ALL stations (max distance): compute on the full set
    from ska_sdp_screen_fitting_b200.kl_screen import station_orders
    orders_full = station_orders(pos, 0, 6)
    scr = pc.run_engine_fit(emu_engine, val, wt, np.asarray(sol.source_positions), pos[sel_a])
    want = gold["screen"][:, sel_f][:, :, sel_a]
    same_order = [i for i, a in enumerate(sel_a) if station_orders(pos[sel_a], 0, 6)[i] == orders_full[a]]
    assert len(same_order) >= 2
    for i in same_order:
        assert np.abs(scr.vals_ph[:, :, i] - want[:, :, i]).max() < pc.TOL_FIT


def test_fit_at_45_directions_against_the_reference_golden(emu_engine, golden_dir):
    """Triggered 5-sigma clip, flagged-subset decompositions, moving orders, non-unit weights:
    the kernels' fit against the reference's own outputs (the cube part runs on the GPU)."""
    pc.check_phase_golden(emu_engine, golden_dir, "phase45", cube=False)


def test_fit_at_100_directions_against_the_reference_golden(emu_engine, golden_dir):
    """The 128-thread fit (above 64 directions: four warps in the vote that compacts the unflagged
    directions, 100-long matrix-vector loops, decompose_kernel<512>) against the reference's own
    outputs at the SKA-Low direction count."""
    pc.check_phase_golden(emu_engine, golden_dir, "phase100", cube=False)


def test_fit_amplitude_type(emu_engine):
    pc.check_fit_amplitude(emu_engine)


def test_fit_tec_type_against_reference_operator_golden(emu_engine, golden_dir):
    pc.check_fit_tec_golden(emu_engine, golden_dir)


def test_tec_soltab_fit_and_dtec_cube_against_reference_golden(emu_engine, golden_dir, tmp_path):
    pc.check_tec_run_golden(emu_engine, golden_dir, tmp_path)


def test_kl_gain_fit_against_reference_golden(emu_engine, golden_dir, repo_solutions):
    pc.check_kl_gain_golden(emu_engine, golden_dir, repo_solutions, full=False)


def test_kl_gain_eval(emu_engine):
    pc.check_kl_gain_eval(emu_engine)
    pc.check_kl_gain_eval(emu_engine, n=11, n_d=5, ncol=3, seed=8)      # ragged: npix % 4 != 0


def test_kl_eval(emu_engine):
    pc.check_kl_eval(emu_engine, n=12, ncol=9)
    pc.check_kl_eval(emu_engine, n=11, n_d=5, ncol=3, seed=8)      # ragged: npix % 4 != 0


def test_kl_eval_tiled_kernel(emu_engine):
    """Large enough (>= 1024 pixels, >= 4 columns) to take the shared-memory tiled kernel:
    4 pixels/thread (ndir <= 52) and 2 pixels/thread (ndir > 52) variants, ragged column tiles."""
    pc.check_kl_eval(emu_engine, n=32, n_d=9, ncol=21, seed=4)
    pc.check_kl_eval(emu_engine, n=36, n_d=60, ncol=19, seed=5)


def test_voronoi_assign_and_fill(emu_engine):
    pc.check_voronoi(emu_engine)
    pc.check_voronoi(emu_engine, n=9, n_d=3, ncol=2, seed=12)      # ragged


def test_voronoi_gain_fill(emu_engine):
    pc.check_voronoi_gain(emu_engine)
    pc.check_voronoi_gain(emu_engine, n=9, n_d=3, ncol=2, seed=12)     # ragged (scalar stores)


def test_voronoi_gain_against_reference_golden(emu_engine, golden_dir, repo_solutions, tmp_path):
    pc.check_voronoi_gain_golden(emu_engine, golden_dir, repo_solutions, tmp_path, full=False)


def test_time_gaps_give_one_file_per_chunk_like_the_reference(emu_engine, golden_dir, repo_solutions, tmp_path):
    pc.check_time_gaps_golden(emu_engine, golden_dir, repo_solutions, tmp_path)


def test_streamed_write_equals_eager_write(emu_engine, tmp_path):
    pc.check_streamed_write(emu_engine, tmp_path)


def test_soltab_write_back(emu_engine, golden_dir, tmp_path):
    pc.check_soltab_write_back(emu_engine, golden_dir, tmp_path)


def test_smoothing(emu_engine):
    pc.check_smooth(emu_engine, shape=(3, 13, 21), sigmas=(0.5, 2.0, 3.0))


def test_smoothed_voronoi_both_forms(emu_engine):
    pc.check_smoothed_voronoi(emu_engine, n=16, ncol=3)
    pc.check_smoothed_voronoi(emu_engine, n=32, n_d=7, ncol=10, seed=7)      # tiled contraction
    pc.check_smoothed_voronoi(emu_engine, n=32, n_d=55, ncol=9, seed=8)


def test_smoothed_gain_voronoi_linear_form(emu_engine):
    pc.check_smoothed_voronoi_gain(emu_engine)


def test_bswap_and_nanfix(emu_engine):
    pc.check_bswap_nanfix(emu_engine)


def test_host_entry_points(emu_engine, emu_library):
    """ssf_voronoi_cube_host / ssf_kl_cube_host: HOST buffers in, HOST slab out."""
    import ctypes

    from oracle.port import klfit, screens, wcs as owcs
    from ska_sdp_screen_fitting_b200 import _lib

    lib = _lib.load(emu_library)
    rng = np.random.default_rng(21)
    n, n_d, ncol = 10, 5, 4
    tmpl = screens.nearest_patch_template(rng.uniform(0, n, n_d), rng.uniform(0, n, n_d), n)
    phases = rng.normal(size=(ncol, n_d))
    out = np.zeros((ncol, 4, n, n), np.float32)
    for sigma in (0.0, 1.0):
        _lib.call("ssf_voronoi_cube_host", tmpl.ctypes.data, phases.ctypes.data, n_d, n, n, ncol,
                  ctypes.c_double(sigma), out.ctypes.data, lib=lib)
        want = screens.voronoi_cube(phases.reshape(1, 1, ncol, n_d), tmpl, sigma)[0, 0]
        assert np.abs(out - want).max() < (pc.TOL_SMOOTH if sigma else 1e-30)
    pp, mid_ra, mid_dec = klfit.piercepoints(pc.make_directions(rng, n_d, 1.0))
    x, y = owcs.kl_grid_coords(126.23, 64.5, (n - 0.5) * 0.2, 0.2, mid_ra, mid_dec)
    _, p_o, _ = klfit.calculate_svd(pp)
    white = rng.normal(0, 0.5, (ncol, n_d)) @ p_o.T
    pp2 = np.ascontiguousarray(pp[:, :2])
    _lib.call("ssf_kl_cube_host", x.ctypes.data, y.ctypes.data, n, n, pp2.ctypes.data, n_d,
              white.ctypes.data, ncol, 100.0, 5.0 / 3.0, 1e-3, out.ctypes.data, lib=lib)
    assert np.abs(out - screens.kl_make_matrix(x, y, pp, white)).max() < pc.TOL_KL


def test_fill_kernels_take_several_grid_passes(emu_library):
    """The fill kernels are grid-strided; the parity cases above never fill one grid pass.  With the
    grid capped at one CTA per SM (experiment hook SSF_FILL_CTAS, read once per process, hence the
    subprocess) a 64 x 64 x 41-column slab takes several passes per thread, ragged at the end."""
    import os
    import subprocess
    import sys

    here = os.path.dirname(os.path.abspath(__file__))
    code = (
        "import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
        "from emu_engine import EmuEngine\n"
        "import parity_checks as pc\n"
        "eng = EmuEngine(%r)\n"
        "pc.check_voronoi(eng, n=64, n_d=7, ncol=41, seed=5)\n"
        "pc.check_voronoi_gain(eng, n=64, n_d=7, ncol=41, seed=6)\n"
        "print('ok')\n" % (os.path.dirname(here), here, os.path.join(here, "emu"), emu_library))
    env = dict(os.environ, SSF_FILL_CTAS="1")
    res = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and res.stdout.strip().endswith("ok"), res.stdout + res.stderr


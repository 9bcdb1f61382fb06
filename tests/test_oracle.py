"""The oracle (oracle.port) against the golden fixtures produced by the REFERENCE's own code
(oracle/make_golden.py under oracle.refshim), and -- when /root/reference is present --
directly against the reference modules."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle import refshim
from oracle.port import klfit, screens, template, wcs

STATIONS = (1, 30, 61)
FREQS = (0, 11)


@pytest.fixture(scope="module")
def kl_golden():
    return np.load(os.path.join(GOLDEN, "kl_c2.npz"))


@pytest.fixture(scope="module")
def vor_golden():
    return np.load(os.path.join(GOLDEN, "voronoi_c1.npz"))


def test_piercepoints_and_singular_values(repo_solutions, kl_golden):
    pp, mid_ra, mid_dec = klfit.piercepoints(np.asarray(repo_solutions.source_positions))
    assert np.abs(pp - kl_golden["piercepoints"]).max() < 1e-9
    assert abs(mid_ra - float(kl_golden["midra"])) < 1e-12
    assert abs(mid_dec - float(kl_golden["middec"])) < 1e-12
    c, _, _ = klfit.calculate_svd(pp)
    sv = np.linalg.svd(c, compute_uv=False)
    # SURVEY.md appendix B known answers
    assert np.allclose(sv, [1544.925, 744.865, 522.017, 187.373, 63.341, 18.775, 8.554], atol=2e-3)


def test_port_fit_matches_reference_golden(repo_solutions, kl_golden):
    sol = repo_solutions
    ref = klfit.get_reference_station(sol.weight, 10)
    assert ref == 0
    out = klfit.run_phase(sol.val, sol.weight, np.asarray(sol.source_positions),
                          np.asarray(sol.station_positions), ref, min(20, 6),
                          stations=STATIONS, freqs=FREQS)
    for f in FREQS:
        for a in STATIONS:
            assert np.abs(out["screen"][:, f, a] - kl_golden["screen"][:, f, a]).max() < 1e-12
            assert np.abs(out["resid"][:, f, a] - kl_golden["resid"][:, f, a]).max() < 1e-12
            assert np.array_equal(out["order"][:, f, a], kl_golden["order"][:, f, a])


def test_port_kl_cube_matches_reference_golden(kl_golden):
    x, y = wcs.kl_grid_coords(126.23, 64.50, 3.33, 0.2, float(kl_golden["midra"]),
                              float(kl_golden["middec"]))
    cube = screens.kl_cube(kl_golden["screen"], kl_golden["piercepoints"], x, y,
                           freqs=list(kl_golden["sub_f"]), stations=list(kl_golden["sub_a"]))
    assert np.abs(cube - kl_golden["cube"]).max() < 1e-6


def test_port_voronoi_cube_matches_reference_golden(repo_solutions, vor_golden):
    ph = screens.voronoi_reference_phases(repo_solutions.val, 0)
    args = dict(freqs=list(vor_golden["sub_f"]), stations=list(vor_golden["sub_a"]))
    cube = screens.voronoi_cube(ph, vor_golden["template"], 0.0, **args)
    assert np.array_equal(cube, vor_golden["cube_nosmooth"])
    cube = screens.voronoi_cube(ph, vor_golden["template"], 0.5, **args)
    assert np.array_equal(cube, vor_golden["cube_smooth"])


def test_restated_smoothing_equals_scipy():
    rng = np.random.default_rng(0)
    block = rng.normal(size=(4, 23, 31)).astype(np.float32)
    for sigma in (0.5, 2.0, 7.0):
        a = screens.smooth_block(block, sigma)
        b = screens.smooth_block_restated(block, sigma)
        assert np.abs(a - b).max() < 2e-6


def test_port_template_matches_reference_golden(vor_golden):
    from ska_sdp_screen_fitting_b200 import skymodel

    pos = skymodel.patch_positions(os.path.join(GOLDEN, "skymodel.txt"))
    names = ["Patch_0", "Patch_1", "Patch_10", "Patch_2", "Patch_3", "Patch_4", "Patch_8"]
    ra = [pos[n][0] for n in names]
    de = [pos[n][1] for n in names]
    tmpl, px, py = template.reference_template(ra, de, 126.23, 64.50, 3.33, 0.2)
    assert np.array_equal(tmpl, vor_golden["template"])
    # test-geometry patch pixels
    assert np.allclose(px, [0.80, 13.64, 10.64, 15.07, 1.67, 5.05, 8.48], atol=6e-3)
    assert np.allclose(py, [11.32, 11.16, 7.56, 0.40, 0.40, 2.76, 5.47], atol=6e-3)
    nearest = screens.nearest_patch_template(px, py, 17)
    # the PIL pipeline is not a pure nearest-patch map: a few boundary pixels differ
    assert 0 < int((nearest != tmpl).sum()) <= 8


def test_reference_test_assertions_hold_on_golden_cubes(repo_solutions, kl_golden, vor_golden):
    """tests/test_fit_screens.py:101-128 / :190-215 (the reference's only pins), applied to
    the golden cubes, two-sided."""
    from ska_sdp_screen_fitting_b200 import skymodel

    pos = skymodel.patch_positions(os.path.join(GOLDEN, "skymodel.txt"))
    val = repo_solutions.val - repo_solutions.val[:, :, 0:1, :]
    for cube, tol, key in ((vor_golden["cube_smooth"], 1e-4, vor_golden), (kl_golden["cube"], 1e-1, kl_golden)):
        sub = val[:, list(key["sub_f"])][:, :, list(key["sub_a"])]
        for d, name in enumerate(repo_solutions.dir):
            ra, de = pos[name.strip("[]")]
            x, y = wcs.voronoi_patch_pixels(ra, de, 126.23, 64.50, 17, 0.2)
            col, row = int(round(float(x))), int(round(float(y)))
            if not (0 <= col < 17 and 0 <= row < 17):
                continue
            assert np.abs(cube[:, :, :, 0, row, col] - np.cos(sub[..., d])).max() < tol
            assert np.abs(cube[:, :, :, 1, row, col] - np.sin(sub[..., d])).max() < tol
            assert np.array_equal(cube[:, :, :, 0], cube[:, :, :, 2])


@pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted")
def test_port_against_the_reference_modules_directly(repo_solutions):
    """Pins oracle.port on a seeded synthetic case with flags, against the reference's own
    _process_station / calculate_kl_screen run under the shim."""
    mods = refshim.install()
    ss = mods["stationscreen"]
    rng = np.random.default_rng(11)
    n_d, n_t = 9, 6
    dirs = np.column_stack([np.deg2rad(126.23 + rng.uniform(-2, 2, n_d)),
                            np.deg2rad(64.5 + rng.uniform(-1.5, 1.5, n_d))]).astype(np.float32)
    pp_ref, _, _ = ss._calculate_piercepoints(np.zeros((1, 3)), dirs)
    pp, _, _ = klfit.piercepoints(dirs)
    assert np.abs(pp_ref - pp).max() < 1e-9
    full_ref = ss._calculate_svd(pp_ref, 100, 5.0 / 3.0, n_d)
    ss.FULL_MATRICES = full_ref
    phase = rng.normal(0, 0.5, (n_d, n_t))
    phase[3, 2] += 2.5                                      # an outlier to be flagged
    weights = np.ones((n_d, n_t), np.float32)
    weights[1, 0] = 0.0
    weights[[2, 5], 4] = 0.0
    names = np.array([f"d{i}" for i in range(n_d)])
    scr_r, w_r, res_r, ord_r = ss._process_station(
        phase.copy(), pp_ref, np.full(n_t, 5.0), weights.copy(), "phase", 2, 5.0, True, names,
        full_ref, 5.0 / 3.0, 100)
    scr_p, w_p, res_p, ord_p = klfit.process_station(
        phase.copy(), pp, np.full(n_t, 5.0), weights.copy(), klfit.calculate_svd(pp))
    assert np.abs(scr_r - scr_p).max() < 1e-10
    assert np.abs(res_r - res_p).max() < 1e-10
    assert np.array_equal(w_r, w_p) and np.array_equal(ord_r, ord_p)


def test_fits_header_matches_the_reference_header(repo_solutions, tmp_path):
    """W1: the product's header cards, key by key and in order, against the header the
    reference wrote for the same cube (make_template_image, processing_utils.py:202-289;
    stored in voronoi_c1.npz by oracle/make_golden.py), and read back from a written file."""
    from ska_sdp_screen_fitting_b200 import fits

    g = np.load(os.path.join(GOLDEN, "voronoi_c1.npz"))
    sol = repo_solutions
    shape = (len(sol.time), len(sol.freq), len(sol.ant), 4, 17, 17)
    cards = fits.aterm_cards(shape, 126.23, 64.50, 0.2, sol.freq, sol.time)
    assert [c[0] for c in cards] == list(g["header_keys"])
    for (key, val, *_), want in zip(cards, g["header_vals"]):
        if isinstance(val, bool) or isinstance(val, str):
            assert str(val) == str(want), key
        else:
            assert float(val) == float(want), key
    writer = fits.CubeWriter(str(tmp_path / "h.fits"), (2, 1, 1, 4, 3, 3),
                             fits.aterm_cards((2, 1, 1, 4, 3, 3), 1.0, 2.0, 0.1, [1e8], [0.0, 8.0]))
    writer.write_times(0, np.arange(72, dtype=np.float32).reshape(2, 1, 1, 4, 3, 3))
    writer.close()
    header, data = fits.read(str(tmp_path / "h.fits"))
    assert header["NAXIS"] == 6 and header["CTYPE3"] == "MATRIX" and header["CDELT6"] == 8.0
    assert np.array_equal(np.asarray(data).ravel(), np.arange(72, dtype=np.float32))
    # the dTEC layout of make_template_image (:187-190, :230-238): no MATRIX axis
    tec = dict((c[0], c[1]) for c in fits.aterm_cards((2, 3, 5, 7, 9), 1.0, 2.0, 0.1, [1e8, 1.1e8, 1.3e8],
                                                     [0.0, 8.0], aterm_type="tec"))
    assert tec["NAXIS"] == 5 and [tec[f"NAXIS{i}"] for i in range(1, 6)] == [9, 7, 5, 3, 2]
    assert [tec[f"CTYPE{i}"] for i in range(1, 6)] == ["RA---SIN", "DEC--SIN", "ANTENNA", "FREQ", "TIME"]
    assert tec["CRPIX1"] == 4.5 and tec["CRPIX2"] == 3.5 and tec["CDELT4"] == 1e7


def test_time_chunks_follow_the_reference_gap_rule():
    """W1 chunking (screen.py:283-294): a new file wherever dt > 1.2 x the smallest dt; two or
    fewer time slots are never split."""
    from ska_sdp_screen_fitting_b200.screen import time_chunks

    t = 8.0 * np.arange(10)
    assert time_chunks(t) == [10]
    gapped = np.concatenate([t[:4], t[4:7] + 100.0, t[7:] + 300.0])
    assert time_chunks(gapped) == [4, 7, 10]
    assert time_chunks(np.array([0.0, 8.0, 17.5])) == [3]            # 9.5 < 1.2 * 8
    assert time_chunks(np.array([0.0, 8.0, 17.7])) == [2, 3]         # 9.7 > 9.6
    assert time_chunks(np.array([0.0, 50.0])) == [2] and time_chunks(np.array([3.0])) == [1]


def test_port_tec_type_matches_reference_operator_golden():
    """_process_station(screen_type="tec") of the reference (golden_tec) against the port."""
    g = np.load(os.path.join(GOLDEN, "tec_station.npz"))
    pp, _, _ = klfit.piercepoints(g["dirs"])
    full = klfit.calculate_svd(pp)
    n_t = g["val"].shape[2]
    for a in range(g["val"].shape[0]):
        scr, w, res, so = klfit.process_station(g["val"][a].copy(), pp, np.full(n_t, g["station_order"][a]),
                                                g["weights"][a].copy(), full, 2, 5.0, True, screen_type="tec")
        assert np.abs(scr - g["screen"][a]).max() < 1e-12 and np.abs(res - g["resid"][a]).max() < 1e-11
        assert np.array_equal(w, g["weights_out"][a]) and np.array_equal(so, g["order"][a])


# ------------------------------------------------------------------------------ gain fixture
@pytest.fixture(scope="module")
def gain_golden():
    return np.load(os.path.join(GOLDEN, "gain_c1.npz"))


def test_port_amplitude_fit_matches_reference_golden(repo_solutions, gain_golden):
    """stationscreen.run on the amplitude soltab (kl_screen.py:114-126): log10 screens,
    non-log residuals, orders and weights of the reference's amplitude_screen000 soltabs."""
    g = gain_golden
    sub_a = [int(a) for a in g["sub_a"]]
    out = klfit.run_amplitude(g["amp_val"], g["amp_weight"], np.asarray(repo_solutions.source_positions),
                              order=4, stations=sub_a)
    assert np.abs(out["screen"][:, :, sub_a] - g["amp_screen"][:, :, sub_a]).max() < 1e-12
    assert np.abs(out["resid"][:, :, sub_a] - g["amp_resid"][:, :, sub_a]).max() < 1e-12
    assert np.array_equal(out["order"][:, :, sub_a], g["amp_order"][:, :, sub_a])
    assert np.array_equal(out["weights"][:, :, sub_a], g["amp_screen_weights"][:, :, sub_a])


def test_port_gain_cubes_match_reference_golden(repo_solutions, gain_golden, kl_golden, vor_golden):
    """Screen.interpolate + the gain branches of both make_matrix implementations."""
    g, sol = gain_golden, repo_solutions
    sub_f, sub_a = [int(f) for f in g["sub_f"]], [int(a) for a in g["sub_a"]]
    # Voronoi: amplitudes interpolated in log space from the raw soltab (log_amps False)
    amp = screens.interpolate_amplitudes(g["amp_val"], g["amp_time"], g["amp_freq"], sol.time, sol.freq, False)
    phases = screens.voronoi_reference_phases(sol.val, 0)
    for key, sigma in (("cube_tessellated_nosmooth", 0.0), ("cube_tessellated", 0.5)):
        cube = screens.voronoi_gain_cube(phases, amp, vor_golden["template"], sigma, freqs=sub_f, stations=sub_a)
        assert np.abs(cube - g[key]).max() <= (1e-6 if sigma else 0.0), key
    # KL: the log10 amplitude screens are interpolated as they are (log_amps True)
    amp_w = screens.interpolate_amplitudes(g["amp_screen"], g["amp_time"], g["amp_freq"], sol.time, sol.freq, True)
    x, y = wcs.kl_grid_coords(126.23, 64.50, 3.33, 0.2, float(kl_golden["midra"]), float(kl_golden["middec"]))
    pp = kl_golden["piercepoints"]
    for fi, f in enumerate(sub_f):
        for ai, a in enumerate(sub_a):
            data = screens.kl_make_matrix(x, y, pp, kl_golden["screen"][:, f, a], amp_w[:, f, a, :, 0],
                                          amp_w[:, f, a, :, 1])
            want = g["cube_kl"][:, fi, ai]
            assert (np.abs(data.astype(np.float32) - want) <= 2e-6 * np.maximum(1.0, np.abs(want))).all()


def test_product_interpolation_matches_scipy(gain_golden, repo_solutions):
    """Screen.interpolate's restated nearest-neighbour rule against scipy's interp1d, including
    exact midpoints and out-of-range abscissae."""
    from ska_sdp_screen_fitting_b200.screen import _interp_axis

    rng = np.random.default_rng(3)
    x_old = np.array([0.0, 2.0, 4.0, 7.0])
    vals = rng.normal(size=(4, 3, 2))
    x_new = np.array([-5.0, 0.0, 1.0, 1.0000001, 3.0, 5.5, 6.9, 7.0, 30.0])
    want = screens.interpolate_amplitudes(vals[:, None], x_old, np.array([1.0]), x_new, np.array([1.0]), True)
    got = _interp_axis(x_old, vals[:, None], 0, x_new, "nearest")
    assert np.array_equal(got, want)
    g, sol = gain_golden, repo_solutions
    want = screens.interpolate_amplitudes(g["amp_val"], g["amp_time"], g["amp_freq"], sol.time, sol.freq, False)
    got = 10 ** _interp_axis(g["amp_freq"], _interp_axis(g["amp_time"], np.log10(g["amp_val"]), 0, sol.time, "nearest"),
                             1, sol.freq, "nearest")
    assert np.array_equal(got, want)


def test_cli_with_string_bounds_writes_the_reference_cube(emu_library, vor_golden, tmp_path):
    """The console entry point (main.py:13-78 of the reference: same flags, defaults, and the
    "[a;b;c;d]" string form of the bounds, make_aterm_images.py:80-87) run as a subprocess on the
    repo data; kernels from the CPU emulator build.  Unsmoothed tessellated cube: bit-exact
    against the reference's, plus the file list and the template file."""
    import subprocess
    import sys

    from ska_sdp_screen_fitting_b200 import fits

    root = os.path.dirname(os.path.dirname(GOLDEN))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, os.path.join(root, "tests", "emu", "emu_engine.py"), emu_library,
           "ska_sdp_screen_fitting_b200.main", os.path.join(GOLDEN, "solutions.h5"),
           "--screen_type", "tessellated", "--outroot", str(tmp_path / "cli"),
           "--bounds_deg", "[124.565; 66.165; 127.895; 62.835]", "--bounds_mid_deg", "[126.23;64.50]",
           "--skymodel", os.path.join(GOLDEN, "skymodel.txt"), "--padding_fraction", "0",
           "--cellsize_deg", "0.2", "--ncpu", "3"]
    done = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert done.returncode == 0, done.stderr[-2000:]
    assert sorted(os.listdir(tmp_path)) == ["cli.txt", "cli_0.fits", "cli_template.fits"]
    with open(tmp_path / "cli.txt", encoding="utf8") as fh:
        assert [ln.strip() for ln in fh] == [str(tmp_path / "cli_0.fits")]
    _, data = fits.read(str(tmp_path / "cli_0.fits"))
    sub = np.asarray(data)[:, list(vor_golden["sub_f"])][:, :, list(vor_golden["sub_a"])]
    assert np.array_equal(sub, vor_golden["cube_nosmooth"])
    # unknown flags / a missing positional are argparse errors (exit status 2), like the reference
    bad = subprocess.run(cmd[:5] + ["--no_such_flag", "1"], env=env, capture_output=True, text=True, timeout=120)
    assert bad.returncode == 2


def test_console_script_is_registered_like_the_reference(tmp_path):
    """pyproject.toml registers the reference's console command
    (/root/reference/pyproject.toml:35-36) and the entry point it names resolves to ``start``;
    ``--help`` runs through it and lists the reference's flags."""
    import subprocess
    import sys
    import tomllib

    root = os.path.dirname(os.path.dirname(GOLDEN))
    with open(os.path.join(root, "pyproject.toml"), "rb") as fh:
        meta = tomllib.load(fh)
    target = meta["project"]["scripts"]["ska-sdp-screen-fitting"]
    module, func = target.split(":")
    assert (module, func) == ("ska_sdp_screen_fitting_b200.main", "start")
    # what the generated console script does: import the module, call the function
    script = tmp_path / "ska-sdp-screen-fitting"
    script.write_text(f"import sys\nfrom {module} import {func}\nsys.exit({func}())\n")
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    done = subprocess.run([sys.executable, str(script), "--help"], env=env, capture_output=True, text=True, timeout=300)
    assert done.returncode == 0, done.stderr[-2000:]
    for flag in ("h5parmfile", "--soltabname", "--screen_type", "--outroot", "--bounds_deg", "--bounds_mid_deg",
                 "--skymodel", "--solsetname", "--padding_fraction", "--cellsize_deg", "--smooth_deg", "--ncpu"):
        assert flag in done.stdout, flag


def test_cli_default_padding_matches_the_reference(emu_library, tmp_path):
    """Default padding_fraction = 1.4 and a cell size that does not divide the padded width
    (make_aterm_images.py:88-97, :116-118): image size, every header card, the template and the
    unsmoothed cube against the reference's (tests/golden/padding_c1.npz, golden_padding)."""
    import subprocess
    import sys

    from ska_sdp_screen_fitting_b200 import fits

    g = np.load(os.path.join(GOLDEN, "padding_c1.npz"))
    root = os.path.dirname(os.path.dirname(GOLDEN))
    env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, os.path.join(root, "tests", "emu", "emu_engine.py"), emu_library,
           "ska_sdp_screen_fitting_b200.main", os.path.join(GOLDEN, "solutions.h5"),
           "--outroot", str(tmp_path / "pad"), "--bounds_deg", "[127.895;62.835;124.565;66.165]",
           "--bounds_mid_deg", "[126.23;64.50]", "--skymodel", os.path.join(GOLDEN, "skymodel.txt"),
           "--cellsize_deg", "0.13"]
    done = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert done.returncode == 0, done.stderr[-2000:]
    header, data = fits.read(str(tmp_path / "pad_0.fits"))
    assert list(header.keys()) == list(g["header_keys"])
    for key, want in zip(g["header_keys"], g["header_vals"]):
        got = header[key]
        if isinstance(got, (bool, str)):
            assert str(got) == str(want), key
        else:
            # cards carry 16 significant digits, as astropy's (".16G") do
            assert abs(float(got) - float(want)) <= 1e-15 * abs(float(want)), key
    sub = np.asarray(data)[:, list(g["sub_f"])][:, :, list(g["sub_a"])]
    assert sub.shape == g["cube"].shape and np.array_equal(sub, g["cube"])
    _, tmpl = fits.read(str(tmp_path / "pad_template.fits"))
    assert np.array_equal(np.asarray(tmpl)[0, 0, 0, 0], g["template"])


def test_h5_reader_chunked_deflate_shuffle_layout():
    """The chunked read path of the minimal HDF5 reader (v1 chunk B-tree, shuffle + deflate
    pipeline, edge chunks, a filter skipped through the chunk's filter mask) on structures laid
    out by hand from the HDF5 file-format specification -- there is no HDF5 library on this
    image to write a real compressed h5parm, and the repo's solutions.h5 is all-contiguous."""
    import struct
    import types
    import zlib

    from ska_sdp_screen_fitting_b200.utils import h5min

    rng = np.random.default_rng(3)
    data = rng.normal(size=(5, 7))
    chunk = (2, 4)
    buf = bytearray(b"\0" * 64)
    entries = []
    for i0 in range(0, 5, chunk[0]):
        for j0 in range(0, 7, chunk[1]):
            block = np.zeros(chunk)
            part = data[i0:i0 + chunk[0], j0:j0 + chunk[1]]
            block[:part.shape[0], :part.shape[1]] = part
            raw = block.astype("<f8").tobytes()
            skip_deflate = (i0, j0) == (2, 4)                     # one chunk stored shuffled only
            shuffled = np.frombuffer(raw, np.uint8).reshape(-1, 8).T.tobytes()
            stored = shuffled if skip_deflate else zlib.compress(shuffled, 5)
            entries.append((len(stored), 2 if skip_deflate else 0, (i0, j0, 0), len(buf)))
            buf += stored
            buf += b"\0" * (-len(buf) % 8)
    tree = len(buf)
    node = b"TREE" + bytes([1, 0]) + struct.pack("<H", len(entries)) + struct.pack("<QQ", h5min._UNDEF, h5min._UNDEF)
    for size, mask, offs, addr in entries:
        node += struct.pack("<II3Q", size, mask, *offs) + struct.pack("<Q", addr)
    node += struct.pack("<II3Q", 0, 0, 6, 8, 0)                  # closing key
    buf += node
    f = object.__new__(h5min.File)
    f._buf, f._so, f._sl = bytes(buf), 8, 8
    ds = types.SimpleNamespace(shape=(5, 7), dtype=np.dtype("<f8"), _layout=("chunked", tree, chunk + (8,)),
                               _filters=[(2, [8]), (1, [5])])
    assert np.array_equal(f._read_chunked(ds), data)
    ds._layout = ("chunked", h5min._UNDEF, chunk + (8,))          # never written: fill value
    assert np.array_equal(f._read_chunked(ds), np.zeros((5, 7)))
    ds._layout = ("chunked", tree, chunk + (8,))
    ds._filters = [(32001, [])]                                    # e.g. blosc: must fail loudly
    with pytest.raises(h5min.H5FormatError):
        f._read_chunked(ds)


@pytest.mark.parametrize("tag", ["phase45", "phase100"])
def test_port_fit_at_real_direction_counts_against_the_reference(tag):
    """oracle.port pinned where the phase-type 5-sigma clip actually fires (> ~30 directions):
    the reference's make_aterm_image(screen_type="kl") on seeded 45- / 100-direction cases
    (tests/golden/phase45.npz, phase100.npz): flags, re-decomposed subsets, orders, residuals,
    and the evaluated cube at the fixture's pixel sample."""
    import parity_checks as pc

    sol, g = pc.phase_golden_solutions(GOLDEN, tag)
    ref = klfit.get_reference_station(sol.weight, 10)
    out = klfit.run_phase(np.asarray(sol.val), np.asarray(sol.weight).copy(), np.asarray(sol.source_positions),
                          np.asarray(sol.station_positions), ref, min(20, len(sol.dir) - 1))
    assert np.abs(out["pp"] - g["piercepoints"]).max() < 1e-9
    assert np.array_equal(out["weights"], g["screen_weights"])
    assert np.array_equal(out["order"], g["order"])
    assert np.abs(out["screen"] - g["screen"]).max() < 1e-10
    assert np.abs(out["resid"] - g["resid"]).max() < 1e-10
    # orders moved away from the per-station starting order somewhere
    from ska_sdp_screen_fitting_b200.kl_screen import station_orders
    start = np.array(station_orders(np.asarray(sol.station_positions), ref, min(20, len(sol.dir) - 1)), float)
    fitted = g["order"] != 0
    assert (g["order"] != start[None, None, :])[fitted].any()
    x, y = wcs.kl_grid_coords(float(g["rad"]), float(g["dec"]), float(g["width_deg"]), float(g["cellsize_deg"]),
                               out["midra"], out["middec"])
    ix = g["pix"]
    for f, a in ((0, 1), (1, 5)):
        data = screens.kl_make_matrix(x[ix], y[ix], out["pp"], out["screen"][:, f, a, :])
        assert np.abs(data.astype(np.float32) - g["cube_sample"][:, f, a]).max() < 1e-6


def test_memory_driven_file_splits_follow_the_reference_rule():
    """screen.py:296-317: chunks longer than max_ntimes are halved, first oversized chunk first,
    until none is left.  Opt-in here (``Screen.max_times_per_file``); restated inline."""
    from ska_sdp_screen_fitting_b200.screen import time_chunks

    def reference_rule(times, max_ntimes):
        delta = times[1:] - times[:-1]
        gaps_ind = np.append(np.where(delta > np.min(delta) * 1.2)[0] + 1, np.array([len(times)]))
        check_gaps = True
        while check_gaps:
            check_gaps = False
            g_start = 0
            for gnum, g_stop in enumerate(gaps_ind.copy()):
                if g_stop - g_start > max_ntimes:
                    gaps_ind = np.insert(gaps_ind, gnum, np.array([g_start + int((g_stop - g_start) / 2)]))
                    check_gaps = True
                    break
                g_start = g_stop
        return [int(g) for g in gaps_ind]

    times = np.arange(100.0) * 8.0
    times[40:] += 100.0
    times[77:] += 50.0
    assert time_chunks(times) == [40, 77, 100]
    for limit in (1, 3, 7, 16, 50, 1000):
        assert time_chunks(times, limit) == reference_rule(times, limit)

"""
TEST INFRASTRUCTURE (oracle.port): numpy restatement of the screen evaluation side of the
reference: KL screen images, Voronoi template + fill, Gaussian smoothing, NaN fix, cube
chunking.  float64 math, float32 storage exactly where the reference stores float32.
"""
import numpy as np
from scipy import ndimage

from . import wcs
from .klfit import BETA, R_0

# ------------------------------------------------------------------------------ KL eval


def kl_kernel_matrix(x_coord, y_coord, pp, r_0=R_0, beta=BETA):
    """kl_screen.py:444-448: K[i, j, d] = -(((X_i-px_d)^2 + (Y_j-py_d)^2 + 0)/r0^2)^(beta/2)/2."""
    dx2 = (x_coord[:, None] - pp[None, :, 0]) ** 2           # [nx, nd]
    dy2 = (y_coord[:, None] - pp[None, :, 1]) ** 2           # [ny, nd]
    d2 = dx2[:, None, :] + dy2[None, :, :] + pp[None, None, :, 2] ** 2
    return -((d2 / r_0**2) ** (beta / 2.0)) / 2.0


def kl_phase_images(kmat, white_td):
    """kl_screen.py:449 for all time slots: val_phase[i, j, t] = K[i, j, :] . white[t, :]."""
    return np.einsum("ijd,td->ijt", kmat, white_td)


def kl_screen_values(x_coord, y_coord, pp, white_td):
    """The screen itself, [t, ny, nx] float64: what calculate_kl_screen returns
    (kl_screen.py:411-449), before make_matrix takes cos / sin of it.  For dTEC solutions this
    is the content of an aterm_type="tec" cube (processing_utils.py:187-190: [t, f, a, y, x])."""
    return kl_phase_images(kl_kernel_matrix(x_coord, y_coord, pp), white_td).T


def kl_make_matrix(x_coord, y_coord, pp, white_td, amp_xx_td=None, amp_yy_td=None):
    """kl_screen.py:192-380: data[t, 4, ny, nx] float64.  Gain solutions: ``amp_xx_td`` /
    ``amp_yy_td`` are the log10 amplitude screens (white values), evaluated with the same
    kernel and raised to the power of ten (:319-365, :374-378)."""
    kmat = kl_kernel_matrix(x_coord, y_coord, pp)
    val_phase = kl_phase_images(kmat, white_td)
    n_t = white_td.shape[0]
    data = np.zeros((n_t, 4, len(y_coord), len(x_coord)))
    if amp_xx_td is None:
        data[:, 0] = np.cos(val_phase.T)                      # :368-373 (note the .T)
        data[:, 2] = data[:, 0]
        data[:, 1] = np.sin(val_phase.T)
        data[:, 3] = data[:, 1]
        return data
    val_amp_xx = 10 ** kl_phase_images(kmat, amp_xx_td)
    val_amp_yy = 10 ** kl_phase_images(kmat, amp_yy_td)
    data[:, 0] = val_amp_xx.T * np.cos(val_phase.T)
    data[:, 2] = val_amp_yy.T * np.cos(val_phase.T)
    data[:, 1] = val_amp_xx.T * np.sin(val_phase.T)
    data[:, 3] = val_amp_yy.T * np.sin(val_phase.T)
    return data


# ------------------------------------------------------------------------------ Voronoi


def voronoi_reference_phases(val_tfad, ref_ind):
    """voronoi_screen.py:74-79."""
    val = np.array(val_tfad, dtype=np.float64)
    return val - val[:, :, ref_ind:ref_ind + 1, :]


def nearest_patch_template(px, py, n):
    """Definition (i) of SURVEY.md 7.4-1: float64 nearest patch per integer pixel (x, y),
    lowest patch index wins ties.  Returns int32 [ny, nx] of patch index + 1."""
    xs = np.arange(n, dtype=np.float64)
    d2 = (xs[None, :, None] - np.asarray(px)[None, None, :]) ** 2 + (
        xs[:, None, None] - np.asarray(py)[None, None, :]) ** 2
    return (np.argmin(d2, axis=2) + 1).astype(np.int32)


def voronoi_make_matrix(template, phases_td, amp_td=None):
    """voronoi_screen.py:169-216.  Phase-only: amplitudes are ones (:88); gain solutions:
    ``amp_td`` [t, d, 2] = (XX, YY) amplitudes on planes 0-1 / 2-3 (:179-214)."""
    n_t, n_d = phases_td.shape
    data = np.zeros((n_t, 4) + template.shape)
    for d in range(n_d):
        ind = np.where(template == d + 1)
        for t in range(n_t):
            axx, ayy = (1.0, 1.0) if amp_td is None else (amp_td[t, d, 0], amp_td[t, d, 1])
            data[t, 0, ind[0], ind[1]] = axx * np.cos(phases_td[t, d])
            data[t, 2, ind[0], ind[1]] = ayy * np.cos(phases_td[t, d])
            data[t, 1, ind[0], ind[1]] = axx * np.sin(phases_td[t, d])
            data[t, 3, ind[0], ind[1]] = ayy * np.sin(phases_td[t, d])
    return data


def interpolate_amplitudes(vals_amp, times_amp, freqs_amp, times_ph, freqs_ph, log_amps,
                           interp_kind="nearest"):
    """screen.py:108-154 (Screen.interpolate) for vals_amp [t', f', a, d, pol]."""
    from scipy import interpolate as si

    vals_amp = np.asarray(vals_amp)
    n_t, n_f = len(times_ph), len(freqs_ph)
    if len(times_amp) == 1:
        new_shape = list(vals_amp.shape)
        new_shape[0], new_shape[1] = n_t, n_f
        return np.resize(vals_amp, new_shape)
    logvals = vals_amp if log_amps else np.log10(vals_amp)
    if vals_amp.shape[0] != n_t:
        logvals = si.interp1d(times_amp, logvals, axis=0, kind=interp_kind, fill_value="extrapolate")(times_ph)
    if vals_amp.shape[1] != n_f:
        logvals = si.interp1d(freqs_amp, logvals, axis=1, kind=interp_kind, fill_value="extrapolate")(freqs_ph)
    return logvals if log_amps else 10 ** logvals


def voronoi_gain_cube(phases_tfad, amp_tfadp, template, smooth_pix=0.0, freqs=None, stations=None):
    """screen.py:328-383 for Voronoi gain solutions: fill, float32 store, smooth, NaN fix."""
    n_t, n_f, n_a, _ = phases_tfad.shape
    freqs = range(n_f) if freqs is None else freqs
    stations = range(n_a) if stations is None else stations
    cube = np.zeros((n_t, len(freqs), len(stations), 4) + template.shape, np.float32)
    for fi, f in enumerate(freqs):
        for ai, a in enumerate(stations):
            cube[:, fi, ai] = voronoi_make_matrix(template, phases_tfad[:, f, a, :], amp_tfadp[:, f, a, :, :2])
            if smooth_pix > 0:
                for t in range(n_t):
                    cube[t, fi, ai] = smooth_block(cube[t, fi, ai], smooth_pix)
    return nan_fix(cube)


# ---------------------------------------------------------------------------- smoothing


def gaussian_weights(sigma, truncate=4.0):
    """scipy.ndimage._filters._gaussian_kernel1d(order=0) as called from gaussian_filter1d:
    radius = int(truncate*sigma + 0.5); w = exp(-0.5 x^2/sigma^2) / sum."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    w = np.exp(-0.5 / (sigma * sigma) * x**2)
    return w / w.sum(), radius


def smooth_block(block_f32, smooth_pix):
    """screen.py:353-362: gaussian_filter on the float32 [4, ny, nx] block, sigma (0, s, s):
    y pass then x pass, 'reflect' boundary, float32 storage between passes."""
    return ndimage.gaussian_filter(block_f32, sigma=(0, smooth_pix, smooth_pix), order=0)


def smooth_block_restated(block_f32, smooth_pix):
    """The same filter without scipy (documents exactly what the CUDA kernel computes):
    per pass, float64 accumulation of w[k] * x[reflect(i+k)], result rounded to float32."""
    w, r = gaussian_weights(smooth_pix)
    out = np.asarray(block_f32, dtype=np.float32)
    for axis in (1, 2):
        n = out.shape[axis]
        idx = np.arange(-r, n + r)
        # scipy 'reflect' = half-sample symmetric: (d c b a | a b c d | d c b a)
        period = 2 * n
        idx = np.mod(idx, period)
        idx = np.where(idx >= n, period - 1 - idx, idx)
        padded = np.take(out, idx, axis=axis).astype(np.float64)
        acc = np.zeros(out.shape, dtype=np.float64)
        for k in range(2 * r + 1):
            sl = [slice(None)] * 3
            sl[axis] = slice(k, k + n)
            acc += w[k] * padded[tuple(sl)]
        out = acc.astype(np.float32)
    return out


def nan_fix(cube_f32):
    """screen.py:364-378: NaN -> 1.0 (planes 0, 2) / 0.0 (planes 1, 3).  cube [..., 4, y, x]."""
    for p in range(4):
        plane = cube_f32[..., p, :, :]
        plane[np.isnan(plane)] = 0.0 if p % 2 else 1.0
    return cube_f32


# ------------------------------------------------------------------------------ chunking


def time_chunks(times):
    """screen.py:283-294: split where dt > 1.2 * min(dt).  (The memory-driven extra splits of
    :296-317 are machine dependent and deliberately not restated -- SURVEY appendix C.)"""
    times = np.asarray(times)
    if len(times) > 2:
        dt = times[1:] - times[:-1]
        gaps = np.where(dt > dt.min() * 1.2)[0] + 1
        return np.append(gaps, len(times))
    return np.array([len(times)])


def header_deltas(freqs, times):
    """processing_utils.py:249-281: (CDELT freq, CDELT time)."""
    freqs = np.asarray(freqs)
    times = np.asarray(times)
    del_freq = np.min(freqs[1:] - freqs[:-1]) if len(freqs) > 1 else 1e8
    if len(times) > 1:
        d = times[1:] - times[:-1]
        del_time = np.min(d[:-1]) if len(times) > 2 else d[0]
    else:
        del_time = 1.0
    return float(del_freq), float(del_time)


# ---------------------------------------------------------------------------- whole cubes


def kl_cube(white_tfad, pp, x_coord, y_coord, freqs=None, stations=None):
    """screen.py:328-383 for KL: float32 cube [t, f, a, 4, y, x] (subset of f / a allowed)."""
    n_t, n_f, n_a, _ = white_tfad.shape
    freqs = range(n_f) if freqs is None else freqs
    stations = range(n_a) if stations is None else stations
    kmat = kl_kernel_matrix(x_coord, y_coord, pp)
    cube = np.zeros((n_t, len(freqs), len(stations), 4, len(y_coord), len(x_coord)), np.float32)
    for fi, f in enumerate(freqs):
        for ai, a in enumerate(stations):
            ph = kl_phase_images(kmat, white_tfad[:, f, a, :])
            cube[:, fi, ai, 0] = np.cos(ph.T)
            cube[:, fi, ai, 2] = cube[:, fi, ai, 0]
            cube[:, fi, ai, 1] = np.sin(ph.T)
            cube[:, fi, ai, 3] = cube[:, fi, ai, 1]
    return nan_fix(cube)


def voronoi_cube(phases_tfad, template, smooth_pix=0.0, freqs=None, stations=None):
    """screen.py:328-383 for Voronoi (phase-only): fill, float32 store, smooth, NaN fix."""
    n_t, n_f, n_a, n_d = phases_tfad.shape
    freqs = range(n_f) if freqs is None else freqs
    stations = range(n_a) if stations is None else stations
    lut = template.astype(np.int64) - 1
    cube = np.zeros((n_t, len(freqs), len(stations), 4) + template.shape, np.float32)
    for fi, f in enumerate(freqs):
        for ai, a in enumerate(stations):
            ph = phases_tfad[:, f, a, :]
            cosv = np.cos(ph)[:, lut]                       # [t, y, x]
            sinv = np.sin(ph)[:, lut]
            cube[:, fi, ai, 0] = cosv
            cube[:, fi, ai, 2] = cosv
            cube[:, fi, ai, 1] = sinv
            cube[:, fi, ai, 3] = sinv
            if smooth_pix > 0:
                for t in range(n_t):
                    cube[t, fi, ai] = smooth_block(cube[t, fi, ai], smooth_pix)
    return nan_fix(cube)

"""
TEST INFRASTRUCTURE (oracle.port): numpy restatement of the Karhunen-Loeve station-screen fit
(phase type) of the reference's ``stationscreen.py``.  float64 throughout, as the reference.

Each function cites the reference lines it follows.  ``pinv(rcond=1e-3)`` is restated with
the ABSOLUTE cutoff scipy 1.7.3 (the reference's pin) applies: singular values <= 1e-3 are
dropped.  The same rule is used by the CUDA path.
"""
import numpy as np

from . import wcs

PINV_CUT = 1e-3
R_0 = 100.0              # stationscreen.py:1046
BETA = 5.0 / 3.0         # stationscreen.py:862


def pinv_abs(a, cut=PINV_CUT):
    """scipy.linalg.pinv(a, rcond=cut) under scipy 1.7.3: atol=cut, rtol=0."""
    u, s, vh = np.linalg.svd(a, full_matrices=False)
    keep = s > cut
    return (vh[keep].T / s[keep]) @ u[:, keep].T


def structure_matrix(pp_a, pp_b, r_0=R_0, beta=BETA):
    """stationscreen.py:417-426 / :570-571 / kl_screen.py:446-448."""
    d2 = ((pp_a[:, None, :] - pp_b[None, :, :]) ** 2).sum(axis=2)
    return -((d2 / r_0**2) ** (beta / 2.0)) / 2.0


def calculate_svd(pp, r_0=R_0, beta=BETA):
    """stationscreen.py:390-430 (_calculate_svd): C, pinv(C), U."""
    c = structure_matrix(pp, pp, r_0, beta)
    u, _, _ = np.linalg.svd(c)
    return c, pinv_abs(c), u


def piercepoints(source_dirs_rad):
    """stationscreen.py:70-110: pierce points = TAN-projected source xy, z = 0."""
    ra = np.rad2deg(np.asarray(source_dirs_rad)[:, 0])
    de = np.rad2deg(np.asarray(source_dirs_rad)[:, 1])
    xy, mid_ra, mid_dec = wcs.getxy(ra, de)
    pp = np.zeros((len(ra), 3))
    pp[:, 0] = xy[0]
    pp[:, 1] = xy[1]
    return pp, mid_ra, mid_dec


def fit_screen(full, pp, phase, weights, order, r_0=R_0, beta=BETA, screen_type="phase"):
    """stationscreen.py:433-594 (_fit_screen, screen_type "phase", "amplitude" or "tec")."""
    n_all = len(phase)
    unfl = np.where(weights > 0.0)[0]
    n = len(unfl)
    if n == n_all:
        c, pinv_c, u = full                                     # :493-494
    else:
        c, pinv_c, u = calculate_svd(pp[unfl], r_0, beta)        # :497-499
    w = np.diag(weights[unfl].astype(weights.dtype))
    uk = u[:, :order]
    inv_u = pinv_abs(uk.T @ (w @ u)[:, :order])                  # :501-504
    if screen_type == "amplitude":                               # :535-548
        with np.errstate(all="ignore"):
            rr1 = uk.T @ (w @ np.log10(phase[unfl]))
        amp_fit_log = pinv_c @ (uk @ (inv_u @ rr1))
        screen_fit = c @ amp_fit_log
    elif screen_type == "tec":                                   # :549-562: linear, one right-hand side
        rr1 = uk.T @ (w @ phase[unfl])
        screen_fit = c @ (pinv_c @ (uk @ (inv_u @ rr1)))
    else:
        fits = []
        for r in (np.cos(phase[unfl]), np.sin(phase[unfl])):     # :509-528
            rr1 = uk.T @ (w @ r)
            fits.append(pinv_c @ (uk @ (inv_u @ rr1)))
        real_fit, imag_fit = fits
        screen_fit = np.arctan2(c @ imag_fit, c @ real_fit)      # :531-533
    white = pinv_c @ screen_fit                                  # :534 / :548
    if n != n_all:                                               # :565-582
        fit_all = np.zeros(n_all)
        fit_all[unfl] = screen_fit
        flagged = np.where(weights <= 0.0)[0]
        for fi in flagged:
            cval = structure_matrix(pp[fi:fi + 1], pp[unfl], r_0, beta)[0]
            fit_all[fi] = cval @ white
        white_all = full[1] @ fit_all
        resid = phase - (10 ** fit_all if screen_type == "amplitude" else fit_all)     # :576-583
    else:
        white_all = white
        model = c @ white
        resid = phase - (10 ** model if screen_type == "amplitude" else model)        # :584-590
    return white_all, resid


def normalize_phase(phase):
    """processing_utils.py:73-98."""
    out = np.fmod(phase, 2.0 * np.pi)
    nans = np.isnan(out)
    out[nans] = 0
    out[out < -np.pi] += 2.0 * np.pi
    out[out > np.pi] -= 2.0 * np.pi
    out[nans] = np.nan
    return out


def flag_outliers(weights, residual, nsigma, screen_type="phase"):
    """stationscreen.py:303-350: 5-sigma clip -- phase: circular std over directions per time;
    amplitude / tec: ONE weighted rms over all directions and times of the station (:338-346).
    Mutates and returns ``weights`` ([ndir, nt]), as the reference does."""
    if not (weights > 0.0).any():
        return weights
    if screen_type in ("amplitude", "tec"):
        nonfl = weights > 0.0
        with np.errstate(all="ignore"):
            std = np.sqrt(np.average(residual[nonfl] ** 2, weights=weights[nonfl], axis=0))
            weights[np.abs(residual) > nsigma * std] = 0.0
        return weights
    res = normalize_phase(residual)
    res_nan = res.copy()
    res_nan[weights == 0.0] = np.nan
    with np.errstate(all="ignore"):
        # processing_utils.py:124-132 (nancircstd over axis 0)
        r_val = np.hypot(np.nanmean(np.sin(res_nan), axis=0), np.nanmean(np.cos(res_nan), axis=0))
        std = np.sqrt(-2 * np.log(r_val))
        weights[np.abs(res) > nsigma * std] = 0.0
    return weights


def circ_chi2(samples, weights):
    """stationscreen.py:353-387."""
    unfl = weights > 0.0
    if not unfl.any():
        return 0.0
    w = weights[unfl]
    m1, sumw = np.average(np.sin(samples[unfl]) ** 2, weights=w, returned=True)
    m2, sumw = np.average(np.cos(samples[unfl]) ** 2, weights=w, returned=True)
    return (1.0 - np.hypot(m1, m2)) * sumw


def log_residual(values, residual):
    """stationscreen.py:664-668, :732-741: log10(a) - log10(|a - resid|)."""
    with np.errstate(all="ignore"):
        return np.log10(values) - np.log10(np.abs(values - residual))


def process_station(phase, pp, screen_order, station_weights, full, niter=2, nsigma=5.0,
                    adjust_order=True, r_0=R_0, beta=BETA, screen_type="phase"):
    """stationscreen.py:597-782 (_process_station, phase).  ``phase``/``station_weights`` are
    [ndir, nt]; ``screen_order`` [nt] is updated in place.  Quirks kept:
    the flagged weights alias the "previous" weights so the first order iteration of every
    iter > 0 pass skips its fit (:657,:669-673,:694-705)."""
    n_dir, n_t = phase.shape
    screen = np.zeros((n_dir, n_t))
    residual = np.zeros((n_dir, n_t))
    station_order = screen_order[0]
    init_w = station_weights.copy()
    for it in range(niter):
        if it > 0:
            # :661-668: phase and tec flag on the residual itself, amplitude on the log residual
            diff = log_residual(phase, residual) if screen_type == "amplitude" else residual.copy()
            station_weights = flag_outliers(init_w, diff, nsigma, screen_type)
        prev_w = init_w
        prev_redchi2 = 0
        norderiter = 4 if (adjust_order and it > 0) else 1
        for t in range(n_t):
            n_unfl = int((station_weights[:, t] > 0.0).sum())
            if n_unfl == 0:
                continue
            if screen_order[t] > n_unfl - 1:
                screen_order[t] = n_unfl - 1
            hit_upper = hit_lower = hit_upper2 = hit_lower2 = False
            sign = 1.0
            for oi in range(norderiter):
                skip_fit = False
                if it > 0 and np.all(station_weights[:, t] == prev_w[:, t]):
                    if not adjust_order:
                        break
                    if oi == 0:
                        skip_fit = True
                if not np.all(station_weights[:, t] == 0.0) and not skip_fit:
                    scr, res = fit_screen(full, pp, phase[:, t], station_weights[:, t],
                                          int(screen_order[t]), r_0, beta, screen_type)
                    screen[:, t] = scr
                    residual[:, t] = res
                if hit_lower2 or hit_upper2:
                    break
                if adjust_order and it > 0:
                    if screen_type == "phase":
                        redchi2 = circ_chi2(residual[:, t], station_weights[:, t]) / (
                            n_unfl - screen_order[t])
                    elif screen_type == "amplitude":                             # :732-741
                        diff = log_residual(phase[:, t], residual[:, t])
                        redchi2 = np.sum(np.square(diff) * station_weights[:, t]) / (
                            n_unfl - screen_order[t])
                    else:                                                        # tec, :742-746
                        redchi2 = np.sum(np.square(residual[:, t]) * station_weights[:, t]) / (
                            n_unfl - screen_order[t])
                    if oi > 0:
                        if redchi2 > 1.0 and prev_redchi2 < redchi2:
                            sign *= -1
                        if redchi2 < 1.0 and prev_redchi2 > redchi2:
                            sign *= -1
                    prev_redchi2 = redchi2
                    order_factor = (n_unfl - screen_order[t]) ** 0.2
                    target = float(screen_order[t]) - sign * order_factor * (1.0 - redchi2)
                    target = max(station_order, target)
                    target = min(int(round(target)), n_unfl - 1)
                    if target <= 0:
                        target = min(station_order, n_unfl - 1)
                    if target == screen_order[t]:
                        break
                    if target == n_unfl - 1:
                        if hit_upper:
                            hit_upper2 = True
                        hit_upper = True
                    if target == station_order:
                        if hit_lower:
                            hit_lower2 = True
                        hit_lower = True
                    screen_order[t] = target
    return screen, station_weights, residual, screen_order


def station_orders(station_pos, ref_ant, order, min_order=5):
    """stationscreen.py:999-1033 with scale_order=True, scale_dist=None."""
    # Positions are the float32 table rows (h5parm.py:731-748) and _get_ant_dist
    # (stationscreen.py:131-135) therefore runs in float32: keep that arithmetic.
    pos = np.asarray(station_pos, dtype=np.float32)
    n = len(pos)
    if ref_ant == -1:
        return [order] * n
    dist = []
    for i in range(n):
        d = pos[ref_ant] - pos[i]
        dist.append(np.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2))
    scale = max(dist)
    out = []
    for d in dist:
        scaled = np.float32(order) * np.sqrt(d / scale)
        out.append(max(min_order, min(order, int(scaled))))
    return out


def get_reference_station(weights_tfad, max_ind=10):
    """processing_utils.py:538-574; weights [time, freq, ant, dir]."""
    n_ant = weights_tfad.shape[2]
    if max_ind is None or max_ind > n_ant:
        max_ind = n_ant
    sums = np.sum(weights_tfad, axis=(0, 1, 3), dtype=float)
    return int(np.where(sums[:max_ind] == sums[:max_ind].max())[0][0])


def run_phase(val_tfad, weight_tfad, source_dirs_rad, station_pos, ref_ant, order,
              niter=2, nsigma=5.0, min_order=5, adjust_order=True, stations=None, freqs=None):
    """stationscreen.py:858-1161 (run) for a scalar phase soltab with axes
    [time, freq, ant, dir].  Returns dict(screen, weights, resid, order [t,f,a], pp, midra,
    middec).  ``stations``/``freqs`` restrict the work to a subset (others stay zero)."""
    val = np.array(val_tfad, dtype=np.float64)
    n_t, n_f, n_a, n_d = val.shape
    wts = np.array(weight_tfad)
    if ref_ant != -1:                                             # :994-997
        val = val - val[:, :, ref_ant:ref_ant + 1, :]
    st_order = station_orders(station_pos, ref_ant, order, min_order)
    pp, mid_ra, mid_dec = piercepoints(source_dirs_rad)
    full = calculate_svd(pp)
    screen = np.zeros((n_t, n_f, n_a, n_d))
    resid = np.zeros((n_t, n_f, n_a, n_d))
    # _process_single_freq returns zeros for stations it skips (stationscreen.py:805,848)
    order_out = np.zeros((n_t, n_f, n_a))
    w_out = wts.copy()
    for f in (range(n_f) if freqs is None else freqs):
        for a in (range(n_a) if stations is None else stations):
            if a == ref_ant:                                      # :818-820
                continue
            ph = val[:, f, a, :].T                                # [dir, time]
            sw = wts[:, f, a, :].T
            if np.all(np.isnan(ph)) or np.all(sw == 0):           # :821-825
                continue
            so = np.full(n_t, float(st_order[a]))                 # :1040-1045
            scr, wgt, res, so = process_station(ph, pp, so, sw, full, niter, nsigma,
                                                adjust_order)
            screen[:, f, a, :] = scr.T
            resid[:, f, a, :] = res.T
            w_out[:, f, a, :] = wgt.T
            order_out[:, f, a] = so
    return dict(screen=screen, weights=w_out, resid=resid, order=order_out, pp=pp,
                midra=mid_ra, middec=mid_dec, full=full, station_order=st_order)


def run_amplitude(val_tfadp, weight_tfadp, source_dirs_rad, order, niter=3, nsigma=5.0,
                  adjust_order=True, stations=None, freqs=None):
    """stationscreen.py:858-1161 (run) for an amplitude soltab [time, freq, ant, dir, pol] as
    KLScreen.fit calls it (kl_screen.py:114-126): niter=3, scale_order=False, ref_ant=-1 (no
    referencing, no station skipped for being the reference).  The screens are log10 values,
    the residuals are not (:1090)."""
    val = np.array(val_tfadp, dtype=np.float64)
    n_t, n_f, n_a, n_d, n_p = val.shape
    wts = np.array(weight_tfadp)
    pp, mid_ra, mid_dec = piercepoints(source_dirs_rad)
    full = calculate_svd(pp)
    screen = np.zeros(val.shape)
    resid = np.zeros(val.shape)
    order_out = np.zeros((n_t, n_f, n_a, n_p))
    w_out = wts.copy()
    for f in (range(n_f) if freqs is None else freqs):
        for p in range(n_p):
            for a in (range(n_a) if stations is None else stations):
                arr = val[:, f, a, :, p].T
                sw = wts[:, f, a, :, p].T.copy()
                if np.all(np.isnan(arr)) or np.all(sw == 0):
                    continue
                so = np.full(n_t, float(order))
                scr, wgt, res, so = process_station(arr, pp, so, sw, full, niter, nsigma, adjust_order,
                                                    screen_type="amplitude")
                screen[:, f, a, :, p] = scr.T
                resid[:, f, a, :, p] = res.T
                w_out[:, f, a, :, p] = wgt.T
                order_out[:, f, a, p] = so
    return dict(screen=screen, weights=w_out, resid=resid, order=order_out, pp=pp, full=full)

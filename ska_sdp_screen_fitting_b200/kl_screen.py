"""
Karhunen-Loeve screens -- drop-in for the reference's ``kl_screen.py:22-449`` with

* ``fit()``: the ``stationscreen.run`` operation (stationscreen.py:858-1161) as a batched
  FP64 GPU fit over all (time, freq, station) systems at once (``Engine.kl_fit``), and
* ``make_matrix()`` / ``write()``: ``calculate_kl_screen`` (kl_screen.py:411-449) as a
  pixel-basis contraction with a fused cos/sin epilogue (``ssf_kl_eval``).

Unlike the reference, the fit results stay in memory: nothing is appended to the input
h5parm (the reference's ``remove_soltabs`` never succeeds, so a second run would read a stale
table -- SURVEY.md section 5).
"""
import queue
import threading
import time

import numpy as np
import torch

from . import wcs
from .engine import BETA, PINV_CUT, R_0, PatternCache
from .screen import Screen


def reference_station(weights, max_ind=10):
    """processing_utils.py:538-574 on weights [time, freq, ant, dir(, pol)]: the first of
    the ``max_ind`` leading antennas with the largest weight sum."""
    n_ant = weights.shape[2]
    if max_ind is None or max_ind > n_ant:
        max_ind = n_ant
    axes = tuple(i for i in range(weights.ndim) if i != 2)
    sums = np.sum(weights, axis=axes, dtype=float)[:max_ind]
    return int(np.nonzero(sums == sums.max())[0][0])


def station_orders(station_positions, ref_ant, order, min_order=5):
    """Per-station screen order from the distance to the reference station
    (stationscreen.py:999-1033, scale_order=True, scale_dist=None).  Positions are the
    float32 rows of the antenna table and the reference does this arithmetic in float32."""
    pos = np.asarray(station_positions, np.float32)
    if ref_ant == -1:
        return [order] * len(pos)
    diff = pos[ref_ant][None, :] - pos
    dist = [np.sqrt(d[0] ** 2 + d[1] ** 2 + d[2] ** 2) for d in diff]
    scale = max(dist)
    return [max(min_order, min(order, int(np.float32(order) * np.sqrt(d / scale)))) for d in dist]


def amplitude_order(n_dir):
    """Order of the amplitude screens (kl_screen.py:95-97)."""
    return min(12, max(3, int(np.round(n_dir / 2))))


class KLScreen(Screen):
    """Class for KL (Karhunen-Lo`eve) screens (reference: kl_screen.py:22-59)."""

    def __init__(self, name, h5parm_filename, skymodel_filename, rad, dec, width_ra, width_dec,
                 solset_name="sol000", phase_soltab_name="phase000", amplitude_soltab_name=None):
        super().__init__(name, h5parm_filename, skymodel_filename, rad, dec, width_ra, width_dec,
                         solset_name=solset_name, phase_soltab_name=phase_soltab_name,
                         amplitude_soltab_name=amplitude_soltab_name)
        self.height = 0.0
        self.beta_val = BETA
        self.r_0 = R_0
        self.piercepoints = None
        self.mid_ra = None
        self.mid_dec = None
        self.screen_order = None          # [t, f, a] final orders (reference: resid soltab weights)
        self.residuals = None             # [t, f, a, d]
        self.weights_out = None           # [t, f, a, d] weights after outlier flagging
        self.eval_variant = -1            # ssf_kl_eval variant (-1 auto, 0 SIMT, 1 tcgen05)
        self._basis = None
        self._basis_key = None
        self._c_full = None
        self._pp_dev = None
        self._so_dev = None
        self.ref_ind = None
        self.station_order = None
        self._pinv_full = None
        self._fit_stream = None
        self._patterns = None
        self.fit_ahead = 2                # fit groups the streaming fit runs ahead of the evaluation
        self.fit_group = 4                # slabs fitted per fit call (amortises the per-call host work)
        self._settle = None

    # ------------------------------------------------------------------ fit
    def setup(self, sol, antenna_weight_sums=None):
        """Once-per-run part of stationscreen.run (stationscreen.py:967-1053): reference
        station, per-station orders, pierce points, and the decomposition of the full
        structure matrix.  ``antenna_weight_sums`` lets a multi-GPU caller pass the
        all-reduced per-antenna weight sums (the reference-station choice is global over
        time, processing_utils.py:560-572)."""
        eng = self.engine
        n_d = len(sol.dir)
        if n_d > 112:
            raise ValueError(f"{n_d} directions exceed the engine limit of 112")
        if antenna_weight_sums is None:
            self.ref_ind = reference_station(sol.weight, 10)              # kl_screen.py:93
        else:
            sums = np.asarray(antenna_weight_sums, float)[:10]
            self.ref_ind = int(np.nonzero(sums == sums.max())[0][0])
        screen_order = min(20, n_d - 1)                                   # kl_screen.py:101
        self.station_order = station_orders(sol.station_positions, self.ref_ind, screen_order)
        # pierce points: TAN-projected float32 source table (stationscreen.py:99-108)
        src = np.asarray(sol.source_positions, np.float32)
        ra_deg, dec_deg = np.rad2deg(src[:, 0]), np.rad2deg(src[:, 1])
        px, py, self.mid_ra, self.mid_dec = wcs.refined_screen_xy(ra_deg, dec_deg)
        self.piercepoints = np.column_stack([px, py, np.zeros(n_d)])
        self._pp_dev = eng.tensor(self.piercepoints[:, :2], torch.float64)
        full = torch.tensor([eng.full_mask(n_d)], dtype=torch.int64, device=eng.device)
        c, pinv, _, _ = eng.kl_decompose(self._pp_dev, full, self.r_0, self.beta_val, PINV_CUT)
        self._c_full, self._pinv_full = c[0].contiguous(), pinv[0].contiguous()
        self._patterns = PatternCache(eng, self._pp_dev, self.r_0, self.beta_val, PINV_CUT)
        self._so_dev = torch.tensor(self.station_order, dtype=torch.float64, device=eng.device)
        self.times_ph = sol.time
        self.freqs_ph = sol.freq
        self.source_names = sol.dir
        self.source_dict = sol.source_dict
        self.source_positions = sol.source_positions
        self.station_names = sol.ant
        self.station_dict = sol.station_dict
        self.station_positions = sol.station_positions

    def fit_device(self, val, weight, check=True):
        """The per-system part of stationscreen.run for a device slab val [t, f, a, d] f64 /
        weight [t, f, a, d] f32 (any number of time slots).  Returns device tensors
        (white [t,f,a,d], resid [t,f,a,d], order [t,f,a], weights_out [t,f,a,d])."""
        eng = self.engine
        t_host = time.perf_counter()
        n_t, n_f, n_a, n_d = val.shape
        ref_ind = self.ref_ind
        if ref_ind != -1:                                                 # stationscreen.py:994-997
            val = val - val[:, :, ref_ind:ref_ind + 1, :]
        # stations skipped by _process_single_freq (stationscreen.py:818-825)
        skip = torch.isnan(val).all(dim=3).all(dim=0) | (weight == 0).all(dim=3).all(dim=0)  # [f, a]
        skip[:, ref_ind] = True
        work_w = torch.where(skip[None, :, :, None], torch.zeros_like(weight), weight).contiguous()
        so = self._so_dev[None, None, :].expand(n_t, n_f, n_a).reshape(-1).contiguous()
        eng.host_tick("fit_prelude", t_host)
        white, resid, order = eng.kl_fit(val.reshape(-1, n_d).contiguous(), work_w.view(-1, n_d),
                                         so, self._pp_dev, niter=2, nsigma=5.0, adjust_order=True,
                                         r0=self.r_0, beta=self.beta_val, cut=PINV_CUT,
                                         cache=self._patterns, check=check)
        t_host = time.perf_counter()
        order = order.view(n_t, n_f, n_a)
        order = torch.where(skip[None], torch.zeros_like(order), order)
        weights_out = torch.where(skip[None, :, :, None], weight, work_w)
        eng.host_tick("fit_post", t_host)
        return white.view(n_t, n_f, n_a, n_d), resid.view(n_t, n_f, n_a, n_d), order, weights_out

    def fit_amplitude_device(self, val, weight, check=True):
        """stationscreen.run on an amplitude soltab as KLScreen.fit calls it (kl_screen.py:114-126:
        niter=3, scale_order=False, no reference station) for device tensors val
        [t, f, a, d, pol] f64 / weight f32.  Returns (white = log10 screens, resid, order
        [t, f, a, pol], weights_out), in the input layout.  The outlier rms of the amplitude
        type spans all time slots of a (freq, station, pol) (stationscreen.py:341-346), so this
        takes the whole (slow) amplitude table at once."""
        eng = self.engine
        n_t, n_f, n_a, n_d, n_p = val.shape
        v = val.permute(0, 1, 2, 4, 3).contiguous()                        # systems [t, f, a, pol] x dir
        w = weight.permute(0, 1, 2, 4, 3).contiguous()
        # (freq, station, pol) groups skipped by _process_single_freq (stationscreen.py:821-825)
        skip = torch.isnan(v).all(dim=4).all(dim=0) | (w == 0).all(dim=4).all(dim=0)
        work_w = torch.where(skip[None, ..., None], torch.zeros_like(w), w).contiguous()
        groups = n_f * n_a * n_p
        so = torch.full((n_t * groups,), float(amplitude_order(n_d)), dtype=torch.float64, device=eng.device)
        white, resid, order = eng.kl_fit(v.view(-1, n_d), work_w.view(-1, n_d), so, self._pp_dev, niter=3,
                                         nsigma=5.0, adjust_order=True, r0=self.r_0, beta=self.beta_val,
                                         cut=PINV_CUT, cache=self._patterns, check=check,
                                         screen_type="amplitude", ngroups=groups)
        shape = (n_t, n_f, n_a, n_p, n_d)
        order = order.view(n_t, n_f, n_a, n_p)
        order = torch.where(skip[None], torch.zeros_like(order), order)
        weights_out = torch.where(skip[None, ..., None], w, work_w)
        back = (0, 1, 2, 4, 3)
        return (white.view(shape).permute(back).contiguous(), resid.view(shape).permute(back).contiguous(),
                order, weights_out.permute(back).contiguous())

    def stream_slabs(self, val_dev, wt_dev, slots, ring, cellsize_deg, first=0, count=None):
        """Streaming form of fit + evaluation for slab-at-a-time processing: yields
        ``(k, slab)`` for k = first, first+1, ... where slab = ring[k % len(ring)] holds the
        cube of time slots [k*slots, (k+1)*slots).  The FP64 fit of slab k+1 runs on a side
        stream while the evaluation kernel of slab k writes the cube (the fit is latency-
        bound and uses a fraction of the SMs; the evaluation is HBM-bound)."""
        eng = self.engine
        n_d = val_dev.shape[-1]
        total = val_dev.shape[0] // slots
        last = total if count is None else min(total, first + count)
        if eng.device.type != "cuda":
            for k in range(first, last):
                white = self.fit_device(val_dev[k * slots:(k + 1) * slots],
                                        wt_dev[k * slots:(k + 1) * slots], check=False)[0]
                slab = ring[k % len(ring)]
                self._fill_columns(white.view(-1, n_d), slab, cellsize_deg, None, 0.0, True)
                yield k, slab
            return
        main = torch.cuda.current_stream(eng.device)
        if self._fit_stream is None:
            self._fit_stream = torch.cuda.Stream(eng.device, priority=-1)   # the fit fills freed SM slots first
        side = self._fit_stream

        side.wait_stream(main)                           # inputs / setup are ready

        group = max(1, int(self.fit_group))
        state = {"last_done": None, "blocked": False}

        def launch_fit(k0, k1):
            """One fit call over the time slots of slabs [k0, k1)."""
            t_host = time.perf_counter()
            with torch.cuda.stream(side):
                white = self.fit_device(val_dev[k0 * slots:k1 * slots],
                                        wt_dev[k0 * slots:k1 * slots], check=False)[0]
                done = torch.cuda.Event()
                done.record(side)
            state["last_done"] = done
            white.record_stream(main)
            self.timings["fit_host_s"] = self.timings.get("fit_host_s", 0.0) + time.perf_counter() - t_host
            self.timings["fit_calls"] = self.timings.get("fit_calls", 0) + 1
            self.timings["fit_slabs"] = self.timings.get("fit_slabs", 0) + (k1 - k0)
            return k0, k1, white, done

        # The fits are issued by a helper thread that runs up to `ahead` groups of `group` slabs
        # in front of the evaluation: the host-side part of a fit (pattern lookup: one small
        # blocking D2H per pass, a fixed cost per call whatever the number of systems) is paid
        # once per group and never keeps the evaluation stream waiting for its next launch.
        ahead = max(1, int(self.fit_ahead))
        ready = queue.Queue(maxsize=ahead)
        stop = threading.Event()

        def issue_fits():
            try:
                torch.cuda.set_device(eng.device)
                for k0 in range(first, last, group):
                    item = launch_fit(k0, min(last, k0 + group))
                    while not stop.is_set():
                        try:
                            ready.put_nowait(item)
                            break
                        except queue.Full:
                            state["blocked"] = True       # as far ahead as allowed
                            time.sleep(0.0002)
                    state["blocked"] = False
                    if stop.is_set():
                        return
            except BaseException as exc:  # surfaced in the consumer
                ready.put(exc)

        def settle():
            """Blocks until the helper has issued every fit it may run ahead (or all of them) and
            orders the main stream after the last one: benchmarks call this on both sides of a
            timed region so that the region contains as many slab fits as slab evaluations."""
            time.sleep(0.001)            # let the helper notice a slot the consumer just freed
            while helper.is_alive() and not state["blocked"]:
                time.sleep(0.0002)
            if state["last_done"] is not None:
                main.wait_event(state["last_done"])

        self._settle = settle

        helper = threading.Thread(target=issue_fits, name="ssf-fit", daemon=True)
        helper.start()
        try:
            k = first
            while k < last:
                item = ready.get()
                if isinstance(item, BaseException):
                    raise item
                k0, k1, white, done = item
                main.wait_event(done)
                white = white.view(k1 - k0, -1, n_d)
                for k in range(k0, k1):
                    slab = ring[k % len(ring)]
                    self._fill_columns(white[k - k0], slab, cellsize_deg, None, 0.0, True)
                    yield k, slab
                k = k1
        finally:
            stop.set()
            helper.join(timeout=30)
            self._settle = None

    def settle_fit_lead(self):
        """See ``stream_slabs``: no-op unless a streamed run is active."""
        if self._settle is not None:
            self._settle()

    def fit(self):
        """kl_screen.py:61-155 + stationscreen.run (phase type)."""
        sol = self._load_solutions(self.input_phase_soltab_name)
        if sol.val.ndim != 4:
            raise NotImplementedError("phase soltabs with a pol axis are not supported (scalarphase only, "
                                      "as the reference's make_matrix assumes: kl_screen.py:263-271)")
        eng = self.engine
        self.setup(sol)
        white, resid, order, weights_out = self.fit_device(eng.tensor(sol.val, torch.float64),
                                                           eng.tensor(sol.weight, torch.float32))
        self._dev_vals = white
        self.vals_ph = white.cpu().numpy()
        self.residuals = resid.cpu().numpy()
        self.screen_order = order.cpu().numpy()
        self.weights_out = weights_out.cpu().numpy()
        if not self.phase_only:                                           # kl_screen.py:113-126, :135-139
            amp = self._load_solutions(self.input_amplitude_soltab_name)
            if amp.val.ndim != 5 or amp.val.shape[-1] < 2:
                raise ValueError("amplitude soltab must be [time, freq, ant, dir, pol] with XX and YY")
            a_white, a_resid, a_order, a_w = self.fit_amplitude_device(eng.tensor(amp.val, torch.float64),
                                                                       eng.tensor(amp.weight, torch.float32))
            self.log_amps = True
            self.vals_amp = a_white.cpu().numpy()
            self.times_amp = amp.time
            self.freqs_amp = amp.freq
            self.amp_residuals = a_resid.cpu().numpy()
            self.amp_screen_order = a_order.cpu().numpy()
            self.amp_weights_out = a_w.cpu().numpy()

    # ------------------------------------------------------------------ evaluation
    def _prepare(self, cellsize_deg, out_dir):
        key = (float(cellsize_deg), float(self.width_ra), float(self.rad), float(self.dec))
        if self._basis is not None and self._basis_key == key:
            return
        eng = self.engine
        x_coord, y_coord = wcs.kl_grid(self.rad, self.dec, self.width_ra, cellsize_deg,
                                       self.mid_ra, self.mid_dec)
        self._basis = eng.kl_basis(eng.tensor(x_coord, torch.float64), eng.tensor(y_coord, torch.float64),
                                   self._pp_dev, self._pinv_full, self.r_0, self.beta_val)
        self._basis_key = key

    def _fill_columns(self, cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan, amp=None):
        # smooth_pix is forced to 0 for KL screens by make_aterm_image (make_aterm_images.py:101-103);
        # NaN -> (1, 0) is fused into the store of ssf_kl_eval either way.
        self._prepare(cellsize_deg, out_dir)
        eng = self.engine
        if amp is not None:
            # gain solutions: phase, log10 XX and log10 YY amplitude screens share the basis
            n_col, n_d = cols.shape
            stacked = torch.stack([cols, amp[..., 0], amp[..., 1]], dim=1).reshape(-1, n_d).contiguous()
            g3 = eng.kl_coeffs(stacked, self._c_full)
            eng.kl_eval_gain(self._basis, g3.view(n_col, 3, -1), n_d, slab, self.eval_variant)
            return
        coeffs = eng.kl_coeffs(cols, self._c_full)
        eng.kl_eval(self._basis, coeffs, cols.shape[1], slab, self.eval_variant)
        if smooth_pix and smooth_pix > 0:
            eng.gauss_smooth(slab, smooth_pix)

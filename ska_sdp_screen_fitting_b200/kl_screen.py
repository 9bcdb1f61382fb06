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

from . import _lib, wcs
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
        self.fit_ahead = 4                # fit blocks the streaming fit runs ahead of the evaluation (slack for host hiccups)
        self.fit_ctas_per_sm = 2          # grid width of the streamed fit (CTAs per SM; the eager fit uses 16)
        self.fit_group = 4                # stream_slabs: slabs fitted per fit call (amortises the per-call host work)
        self.keep_fit_results = True      # streamed write(): bring white / resid / order / weights back to the host
        self.write_soltabs = False        # append the screen soltabs to the input h5parm after the fit (write_back)
        self._settle = None
        self._host_sol = None             # (val, weight) host arrays of this rank's time rows
        self._skip_dev = None
        self._fitted = None               # results of fit_now()
        self._streamed = None             # pinned host copies of the results of a streamed write()

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

    def fit_device(self, val, weight, check=True, skip=None):
        """The per-system part of stationscreen.run for a device slab val [t, f, a, d] f64 /
        weight [t, f, a, d] f32 (any number of time slots).  Returns device tensors
        (white [t,f,a,d], resid [t,f,a,d], order [t,f,a], weights_out [t,f,a,d]).
        ``skip`` [f, a] bool: the (freq, station) pairs _process_single_freq leaves out
        (all-NaN values or all-zero weights over ALL times, stationscreen.py:818-825); when the
        slab is only part of the time axis the caller passes the global answer."""
        eng = self.engine
        t_host = time.perf_counter()
        n_t, n_f, n_a, n_d = val.shape
        ref_ind = self.ref_ind
        if ref_ind != -1:                                                 # stationscreen.py:994-997
            val = val - val[:, :, ref_ind:ref_ind + 1, :]
        if skip is None:
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

    def fit_tec_device(self, val, weight, check=True):
        """stationscreen.run on a soltab of type "tec" (val [t, f, a, d] f64 / weight f32): the
        flow of the phase type -- referencing, per-station orders, niter = 2 -- with the tec
        branches of the operators (values fitted as they are :549-562, rms clip over all times of
        a (freq, station) on the plain residual :338-346, plain chi-square :742-746)."""
        eng = self.engine
        n_t, n_f, n_a, n_d = val.shape
        if self.ref_ind != -1:
            val = val - val[:, :, self.ref_ind:self.ref_ind + 1, :]
        skip = torch.isnan(val).all(dim=3).all(dim=0) | (weight == 0).all(dim=3).all(dim=0)
        skip[:, self.ref_ind] = True
        work_w = torch.where(skip[None, :, :, None], torch.zeros_like(weight), weight).contiguous()
        so = self._so_dev[None, None, :].expand(n_t, n_f, n_a).reshape(-1).contiguous()
        white, resid, order = eng.kl_fit(val.reshape(-1, n_d).contiguous(), work_w.view(-1, n_d), so, self._pp_dev,
                                         niter=2, nsigma=5.0, adjust_order=True, r0=self.r_0, beta=self.beta_val,
                                         cut=PINV_CUT, cache=self._patterns, check=check, screen_type="tec",
                                         ngroups=n_f * n_a)
        order = torch.where(skip[None], torch.zeros_like(order.view(n_t, n_f, n_a)), order.view(n_t, n_f, n_a))
        weights_out = torch.where(skip[None, :, :, None], weight, work_w)
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

    # ------------------------------------------------------------------ streamed fit
    def _fit_blocks(self, fetch, bounds, skip=None):
        """Generator over fitted blocks of time slots, with the fits running ahead of the consumer.

        ``bounds`` = [(r0, r1), ...] row ranges in consumption order; ``fetch(r0, r1)`` returns the
        device tensors (val [rows, f, a, d] f64, weight f32) of a block and is called on the fit
        stream by a helper thread (host staging + H2D copy for host-resident solutions, a slice
        for device-resident ones).  Yields ``(r0, r1, fit)`` with fit = the tuple ``fit_device``
        returns, after ordering the consumer's current stream behind that block's fit.

        The FP64 fit of the next blocks runs on a high-priority side stream while the consumer's
        evaluation kernels write the cube (the fit is latency-bound and uses a fraction of the
        SMs; the evaluation is HBM-bound), up to ``fit_ahead`` blocks in front of it."""
        eng = self.engine
        if eng.device.type != "cuda":
            for r0, r1 in bounds:
                val, wt = fetch(r0, r1)
                yield r0, r1, self.fit_device(val, wt, check=False, skip=skip)
            return
        main = torch.cuda.current_stream(eng.device)
        if self._fit_stream is None:
            self._fit_stream = torch.cuda.Stream(eng.device, priority=-1)   # the fit fills freed SM slots first
        side = self._fit_stream
        side.wait_stream(main)                           # inputs / setup are ready
        # Give the side stream's allocator pool one large cached segment up front: every tensor
        # of a fit call (a few dozen MB per block) is then carved from it, and the steady state
        # never reaches cudaMalloc -- which waits for the work queued on the device (the main
        # thread runs many evaluation launches ahead) and so stalled the fit thread for tens of
        # ms at random (seen as fit_host_ms_per_slab 6-7 ms instead of 0.5 in two bench runs).
        if not getattr(self, "_fit_pool_primed", False):
            rows = max((r1 - r0) for r0, r1 in bounds) if bounds else 0
            n_f, n_a = len(self.freqs_ph), len(self.station_names)
            per_block = rows * n_f * n_a * self._pp_dev.shape[0] * 8
            with torch.cuda.stream(side):
                reserve = torch.empty(int(min(max(24 * per_block, 256 << 20), 4 << 30)), dtype=torch.uint8,
                                      device=eng.device)
                del reserve
            self._fit_pool_primed = True
        state = {"last_done": None, "blocked": False}

        def launch_fit(r0, r1):
            t_host = time.perf_counter()
            with torch.cuda.stream(side):
                val, wt = fetch(r0, r1)
                fit = self.fit_device(val, wt, check=False, skip=skip)
                done = torch.cuda.Event()
                done.record(side)
            state["last_done"] = done
            for t in fit:
                t.record_stream(main)
            self.timings["fit_host_s"] = self.timings.get("fit_host_s", 0.0) + time.perf_counter() - t_host
            self.timings["fit_calls"] = self.timings.get("fit_calls", 0) + 1
            self.timings["fit_rows"] = self.timings.get("fit_rows", 0) + (r1 - r0)
            return r0, r1, fit, done

        # The fits are issued by a helper thread that runs up to `ahead` blocks in front of the
        # evaluation: the host side of a fit call (a few dozen launches) never keeps the
        # evaluation stream waiting for its next launch.
        ahead = max(1, int(self.fit_ahead))
        ready = queue.Queue(maxsize=ahead)
        stop = threading.Event()

        def issue_fits():
            try:
                torch.cuda.set_device(eng.device)
                for r0, r1 in bounds:
                    item = launch_fit(r0, r1)
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
            except BaseException as exc:  # noqa: BLE001 - surfaced in the consumer
                ready.put(exc)

        def settle():
            """Blocks until the helper has issued every fit it may run ahead (or all of them) and
            orders the main stream after the last one: benchmarks call this on both sides of a
            timed region so that the region contains as many block fits as block evaluations."""
            time.sleep(0.001)            # let the helper notice a slot the consumer just freed
            while helper.is_alive() and not state["blocked"]:
                time.sleep(0.0002)
            if state["last_done"] is not None:
                main.wait_event(state["last_done"])

        self._settle = settle
        # a fit that runs beside the cube kernels is launched narrow (ssf_kl_fit_ctas_per_sm): its
        # CTAs then fit next to the resident evaluation CTAs instead of replacing them
        eng._call("ssf_kl_fit_ctas_per_sm", int(self.fit_ctas_per_sm))
        helper = threading.Thread(target=issue_fits, name="ssf-fit", daemon=True)
        helper.start()
        try:
            for _ in bounds:
                item = ready.get()
                if isinstance(item, BaseException):
                    raise item
                r0, r1, fit, done = item
                main.wait_event(done)
                yield r0, r1, fit
        finally:
            stop.set()
            helper.join(timeout=30)
            self._settle = None
            eng._call("ssf_kl_fit_ctas_per_sm", 0)

    def stream_slabs(self, val_dev, wt_dev, slots, ring, cellsize_deg, first=0, count=None, out_flags=0):
        """Fit + evaluation of device-resident solutions slab by slab: yields ``(k, slab)`` for
        k = first, first+1, ... where slab = ring[k % len(ring)] holds the cube of time slots
        [k*slots, (k+1)*slots).  Fits are issued ``fit_group`` slabs at a time (``_fit_blocks``)."""
        n_d = val_dev.shape[-1]
        total = val_dev.shape[0] // slots
        last = total if count is None else min(total, first + count)
        group = max(1, int(self.fit_group))
        bounds = [(k0 * slots, min(last, k0 + group) * slots) for k0 in range(first, last, group)]

        def fetch(r0, r1):
            return val_dev[r0:r1], wt_dev[r0:r1]

        for r0, r1, fit in self._fit_blocks(fetch, bounds):
            white = fit[0].view((r1 - r0) // slots, -1, n_d)
            for j in range(white.shape[0]):
                k = r0 // slots + j
                slab = ring[k % len(ring)]
                self._fill_columns(white[j], slab, cellsize_deg, None, 0.0, True, out_flags=out_flags)
                yield k, slab

    def settle_fit_lead(self):
        """See ``_fit_blocks``: no-op unless a streamed run is active."""
        if self._settle is not None:
            self._settle()

    def fit(self):
        """kl_screen.py:61-155 + stationscreen.run (phase type).

        The per-system fits are DEFERRED: this reads the rank's time rows, fixes everything that is
        global over time (reference station, stations to skip, per-station orders, pierce points,
        the decomposition of the full structure matrix) and returns.  ``write()`` then fits block
        by block on a side stream underneath the evaluation of the previous block; asking for a
        result first (``vals_ph``, ``residuals``, ``screen_order``, ``weights_out``,
        ``make_matrix``) runs the whole fit at once (``fit_now``)."""
        sol, times = self._load_rows(self.input_phase_soltab_name)
        if sol.val.ndim != 4:
            raise NotImplementedError("phase soltabs with a pol axis are not supported (scalarphase only, "
                                      "as the reference's make_matrix assumes: kl_screen.py:263-271)")
        eng = self.engine
        val, weight = np.asarray(sol.val), np.asarray(sol.weight)
        # global over time: per-antenna weight sums (processing_utils.py:560-572) and the
        # (freq, station) pairs _process_single_freq skips (stationscreen.py:818-825)
        sums = np.zeros(val.shape[2])
        all_nan = np.ones(val.shape[1:3], bool)
        all_zero = np.ones(val.shape[1:3], bool)
        for t0 in range(0, val.shape[0], 64):
            sums += np.sum(weight[t0:t0 + 64], axis=(0, 1, 3), dtype=float)
            all_nan &= np.isnan(val[t0:t0 + 64]).all(axis=(0, 3))
            all_zero &= (weight[t0:t0 + 64] == 0).all(axis=(0, 3))
        sums = self._all_reduce(sums, "sum")
        flags = self._all_reduce(np.stack([all_nan, all_zero]).astype(np.int32), "min").astype(bool)
        self.setup(sol, antenna_weight_sums=sums)
        self.times_ph = times
        skip = flags[0] | flags[1]
        skip[:, self.ref_ind] = True
        self._skip_dev = torch.as_tensor(skip).to(eng.device)
        self._host_sol = (val, weight)
        self._fitted = None
        self._streamed = None
        if str(sol.soltype).lower().startswith("tec"):
            # dTEC solutions: the values themselves are fitted (stationscreen.py:549-562) and the
            # outlier rms spans all time slots of a (freq, station) (:338-346), so the fit takes
            # the whole table at once; write() then makes the 5-axis cube of the screens
            if self._all_reduce(np.ones(1), "sum")[0] > 1:
                raise NotImplementedError("tec screens are fitted on one rank (their outlier rms spans all times)")
            if not self.phase_only:
                raise ValueError("a tec soltab cannot be combined with an amplitude soltab")
            self.aterm_type = "tec"
            white, resid, order, w_out = self.fit_tec_device(eng.tensor(val, torch.float64),
                                                             eng.tensor(weight, torch.float32))
            self._fitted = {"dev": white, "white": white.cpu().numpy(), "resid": resid.cpu().numpy(),
                            "order": order.cpu().numpy(), "weights": w_out.cpu().numpy()}
        if not self.phase_only:                                           # kl_screen.py:113-126, :135-139
            amp = self._load_solutions(self.input_amplitude_soltab_name)
            if amp.val.ndim != 5 or amp.val.shape[-1] < 2:
                raise ValueError("amplitude soltab must be [time, freq, ant, dir, pol] with XX and YY")
            # the slow amplitude table is small and its outlier rms spans all times of a (freq,
            # station, pol): every rank fits all of it (replicated, deterministic)
            a_white, a_resid, a_order, a_w = self.fit_amplitude_device(eng.tensor(amp.val, torch.float64),
                                                                       eng.tensor(amp.weight, torch.float32))
            self.log_amps = True
            self.vals_amp = a_white.cpu().numpy()
            self.times_amp = amp.time
            self.freqs_amp = amp.freq
            self._amp_fit = {"white": self.vals_amp}                       # on the slow grid (interpolate() resamples vals_amp)
            self._amp_axes = (np.asarray(amp.time), np.asarray(amp.freq))
            self.amp_residuals = a_resid.cpu().numpy()
            self.amp_screen_order = a_order.cpu().numpy()
            self.amp_weights_out = a_w.cpu().numpy()

    def _fetch_host(self, r0, r1):
        """Rows [r0, r1) of the host-resident solutions -> device (pinned staging, async copy on
        the current stream)."""
        eng = self.engine
        val, weight = self._host_sol
        v = torch.from_numpy(np.ascontiguousarray(val[r0:r1], dtype=np.float64))
        w = torch.from_numpy(np.ascontiguousarray(weight[r0:r1], dtype=np.float32))
        if eng.device.type == "cuda":
            v, w = v.pin_memory(), w.pin_memory()
        self.timings["h2d_bytes"] = self.timings.get("h2d_bytes", 0) + v.numel() * 8 + w.numel() * 4
        return v.to(eng.device, non_blocking=True), w.to(eng.device, non_blocking=True)

    def fit_now(self):
        """Fits all of this rank's time rows at once and keeps the results on the device
        (``_dev_vals``) and as the reference's host attributes."""
        if self._fitted is not None:
            return self._fitted
        if self._streamed is not None:                       # a streamed write() already fitted everything
            self._fitted = {k: v.numpy() for k, v in self._streamed.items()}
            self._fitted["dev"] = None
            return self._fitted
        if self._host_sol is None:
            raise RuntimeError("KLScreen.fit() has not been called")
        eng = self.engine
        n_rows = self._host_sol[0].shape[0]
        parts = []
        rows = max(1, int(self.block_columns * 4 // max(1, len(self.freqs_ph) * len(self.station_names))))
        for r0 in range(0, n_rows, rows):
            v, w = self._fetch_host(r0, min(n_rows, r0 + rows))
            parts.append(self.fit_device(v, w, check=False, skip=self._skip_dev))
        eng.check_fit_status()
        white, resid, order, weights_out = (torch.cat([p[i] for p in parts]) for i in range(4))
        self._fitted = {"dev": white, "white": white.cpu().numpy(), "resid": resid.cpu().numpy(),
                        "order": order.cpu().numpy(), "weights": weights_out.cpu().numpy()}
        return self._fitted

    # ------------------------------------------------------------------ soltab write-back
    def write_back(self, filename=None):
        """The side effect of the reference's fit (stationscreen.py:1087-1159): appends the soltabs
        ``phase_screen000`` (whitened screens, weights after flagging) and ``phase_screen000resid``
        (residuals; the final screen ORDER stored as the weights) to the solset of the h5parm, with
        the attributes ``beta / r_0 / height / midra / middec`` and the ``piercepoint`` table on the
        former -- and ``amplitude_screen000`` / ``...resid`` (pol axis, log10 screens) for gain
        solutions.  Opt-in (``write_soltabs``): the reference always does this, but here the file
        is modified by this package's own HDF5 appender (utils/h5write.py; no HDF5 library on the
        image).  With ``torch.distributed`` the ranks' time rows are gathered on rank 0."""
        from .screen import _dist_info
        from .utils import h5write

        filename = filename or self.input_h5parm_filename
        if filename is None:
            raise ValueError("no h5parm file to write to")
        res = self.fit_now()
        parts = {k: res[k] for k in ("white", "resid", "order", "weights")}
        rank, world, dist = _dist_info()
        if world > 1:
            boxes = [None] * world if rank == 0 else None
            dist.gather_object((self._local_rows(), parts), boxes, dst=0)
            if rank != 0:
                return
            n_t = len(self.times_ph)
            full = {k: np.zeros((n_t,) + v.shape[1:], v.dtype) for k, v in parts.items()}
            for rows, part in boxes:
                for k in full:
                    full[k][rows] = part[k]
            parts = full
        ant = np.array([str(a).encode() for a in self.station_names])
        dirs = np.array([str(d).encode() for d in self.source_names])
        axes = ["time", "freq", "ant", "dir"]
        vals = [np.asarray(self.times_ph, np.float64), np.asarray(self.freqs_ph, np.float64), ant, dirs]
        attrs = {"beta": float(self.beta_val), "r_0": float(self.r_0), "height": 0.0,
                 "midra": float(self.mid_ra), "middec": float(self.mid_dec)}
        solset = self.input_solset_name
        kind = "tec" if self.aterm_type == "tec" else "phase"          # f"{screen_type}screen", stationscreen.py:1097
        h5write.append_soltab(filename, solset, f"{kind}_screen000", f"{kind}screen", axes, vals, parts["white"],
                              parts["weights"], attrs=attrs, extra={"piercepoint": self.piercepoints})
        order = np.broadcast_to(parts["order"][..., None], parts["resid"].shape)
        h5write.append_soltab(filename, solset, f"{kind}_screen000resid", f"{kind}screenresid", axes, vals,
                              parts["resid"], order)
        if not self.phase_only:
            axes_p = axes + ["pol"]
            vals_p = [np.asarray(self._amp_axes[0], np.float64), np.asarray(self._amp_axes[1], np.float64), ant, dirs,
                      np.array([b"XX", b"YY"])]
            h5write.append_soltab(filename, solset, "amplitude_screen000", "amplitudescreen", axes_p, vals_p,
                                  self._amp_fit["white"], self.amp_weights_out, attrs=attrs,
                                  extra={"piercepoint": self.piercepoints})
            a_order = np.broadcast_to(self.amp_screen_order[:, :, :, None, :], self.amp_residuals.shape)
            h5write.append_soltab(filename, solset, "amplitude_screen000resid", "amplitudescreenresid", axes_p,
                                  vals_p, self.amp_residuals, a_order)

    # the reference's result attributes (kl_screen.py:131-154), computed on first use
    @property
    def vals_ph(self):
        return None if self._host_sol is None else self.fit_now()["white"]

    @vals_ph.setter
    def vals_ph(self, value):
        if value is not None:
            self._fitted = dict(self._fitted or {}, white=value)

    @property
    def residuals(self):
        return None if self._host_sol is None else self.fit_now()["resid"]

    @residuals.setter
    def residuals(self, value):
        if value is not None:
            self._fitted = dict(self._fitted or {}, resid=value)

    @property
    def screen_order(self):
        return None if self._host_sol is None else self.fit_now()["order"]

    @screen_order.setter
    def screen_order(self, value):
        if value is not None:
            self._fitted = dict(self._fitted or {}, order=value)

    @property
    def weights_out(self):
        return None if self._host_sol is None else self.fit_now()["weights"]

    @weights_out.setter
    def weights_out(self, value):
        if value is not None:
            self._fitted = dict(self._fitted or {}, weights=value)

    @property
    def _dev_vals(self):
        if self._host_sol is None:
            return None
        fitted = self.fit_now()
        if fitted.get("dev") is None:
            fitted["dev"] = self.engine.tensor(fitted["white"], torch.float64)
        return fitted["dev"]

    @_dev_vals.setter
    def _dev_vals(self, value):
        pass

    def _blocks(self, local0, local1):
        """write()'s source of column blocks.  After an eager fit the base class serves the
        device-resident results; otherwise the blocks are fitted here as they are consumed
        (``_fit_blocks``), the solutions of each block going host -> device right before its fit,
        and -- unless ``keep_fit_results`` is off -- the results coming back to pinned host arrays
        so the reference's attributes exist after the run without a second fit."""
        if self._fitted is not None and self._fitted.get("white") is not None:
            yield from super()._blocks(local0, local1)
            return
        eng = self.engine
        n_f, n_a = len(self.freqs_ph), len(self.station_names)
        per_t = n_f * n_a
        n_d = self._host_sol[0].shape[-1]
        n_rows = self._host_sol[0].shape[0]
        rows = max(1, int(self.block_columns // per_t))
        bounds = [(r0, min(local1, r0 + rows)) for r0 in range(local0, local1, rows)]
        keep = None
        if self.keep_fit_results:
            if self._streamed is None:
                pin = eng.device.type == "cuda"
                self._streamed = {
                    "white": torch.empty((n_rows, n_f, n_a, n_d), dtype=torch.float64, pin_memory=pin),
                    "resid": torch.empty((n_rows, n_f, n_a, n_d), dtype=torch.float64, pin_memory=pin),
                    "order": torch.empty((n_rows, n_f, n_a), dtype=torch.float64, pin_memory=pin),
                    "weights": torch.empty((n_rows, n_f, n_a, n_d), dtype=torch.float32, pin_memory=pin)}
            keep = self._streamed
        for r0, r1, fit in self._fit_blocks(self._fetch_host, bounds, skip=self._skip_dev):
            white = fit[0]
            if keep is not None:
                for name, t in zip(("white", "resid", "order", "weights"), fit):
                    keep[name][r0:r1].copy_(t, non_blocking=True)
            amp = None if self.phase_only else self._dev_amp[r0:r1].reshape(-1, n_d, 2)
            yield r0, r1, white.view(-1, n_d), amp

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

    def _out_format(self, smooth_pix):
        # KL screens are never smoothed (make_aterm_images.py:101-103), so the evaluation kernel's
        # store is the last touch of every value: FITS byte order always, plane pairs when the
        # solutions are phase-only (planes 2, 3 = planes 0, 1, kl_screen.py:368-373)
        if self.aterm_type == "tec":
            return _lib.OUT_BIG_ENDIAN, 1
        if self.phase_only:
            return _lib.OUT_BIG_ENDIAN | _lib.OUT_PLANES2, 2
        return _lib.OUT_BIG_ENDIAN, 4

    def _fill_columns(self, cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan, amp=None, out_flags=0):
        # smooth_pix is forced to 0 for KL screens by make_aterm_image (make_aterm_images.py:101-103);
        # NaN -> (1, 0) is fused into the store of ssf_kl_eval either way.
        self._prepare(cellsize_deg, out_dir)
        eng = self.engine
        if self.aterm_type == "tec" and slab.shape[1] == 1:
            # dTEC cube: the screen values themselves, one plane per column
            eng.kl_eval_tec(self._basis, eng.kl_coeffs(cols.contiguous(), self._c_full), cols.shape[1], slab,
                            self.eval_variant, out_flags & _lib.OUT_BIG_ENDIAN)
            return
        if amp is not None:
            # gain solutions: phase, log10 XX and log10 YY amplitude screens share the basis
            n_col, n_d = cols.shape
            stacked = torch.stack([cols, amp[..., 0], amp[..., 1]], dim=1).reshape(-1, n_d).contiguous()
            g3 = eng.kl_coeffs(stacked, self._c_full)
            eng.kl_eval_gain(self._basis, g3.view(n_col, 3, -1), n_d, slab, self.eval_variant, out_flags)
            return
        coeffs = eng.kl_coeffs(cols.contiguous(), self._c_full)
        eng.kl_eval(self._basis, coeffs, cols.shape[1], slab, self.eval_variant, out_flags)
        if smooth_pix and smooth_pix > 0:
            if out_flags:
                raise ValueError("smoothing needs native 4-plane slabs")
            eng.gauss_smooth(slab, smooth_pix)

    def make_tec_matrix(self, t_start_index, t_stop_index, freq_ind, stat_ind, cellsize_deg):
        """dTEC screens of one (freq, station): float64 [t, ny, nx] -- the screen itself without
        the cos / sin of make_matrix, i.e. the content of an ``aterm_type="tec"`` cube
        (processing_utils.py:187-190) for solutions fitted with the tec screen type."""
        self._prepare(cellsize_deg, None)
        eng = self.engine
        n_y, n_x = self.image_shape(cellsize_deg)
        cols = self._dev_vals[t_start_index:t_stop_index, freq_ind, stat_ind, :].contiguous()
        out = eng.empty((cols.shape[0], n_y, n_x), torch.float32)
        eng.kl_eval_tec(self._basis, eng.kl_coeffs(cols, self._c_full), cols.shape[1], out, self.eval_variant)
        return out.double().cpu().numpy()

"""
Device-side orchestration of the hot path: torch tensors supply device memory and the
current CUDA stream; every computation is a call into ``libssf_b200.so`` (include/ssf.h).

torch is plumbing only (allocation, views, streams); nothing here computes a screen in Python
or falls back to the CPU.
"""
import math
import threading
import time
import weakref

import numpy as np
import torch

from . import _lib

R_0 = 100.0            # stationscreen.py:1046
BETA = 5.0 / 3.0       # stationscreen.py:862
PINV_CUT = 1e-3        # stationscreen.py:427, :504 (absolute cutoff, scipy 1.7.3 semantics)
MAX_DIR = 112          # SSF_MAX_DIR


def _ptr(t):
    return None if t is None else t.data_ptr()


class Engine:
    """Thin, stateless-per-call wrapper of ``libssf_b200.so`` on one CUDA device.  There is no
    other device: without the library or without CUDA the constructor raises."""

    def __init__(self, device=None):
        lib = _lib.load()
        device = torch.device("cuda" if device is None else device)
        if device.type != "cuda":
            raise _lib.SsfLibraryError("libssf_b200.so needs a CUDA device; there is no CPU path")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self._init_state(lib, device)

    def _init_state(self, lib, device):
        self._lib = lib
        self.device = device
        self._fit_status = None
        self.timers = None
        self.host_times = {}
        self._cpu_mark = {}
        self._caches = weakref.WeakSet()

    # ------------------------------------------------------------------ helpers
    def _call(self, name, *args):
        return _lib.call(name, *args, lib=self._lib)

    def _timed_call(self, key, name, *args):
        """Like _call, bracketed by CUDA events on the launching stream when ``self.timers`` is a
        dict (bench.py: per-kernel durations measured live inside the timed region)."""
        if self.timers is None or self.device.type != "cuda":
            return self._call(name, *args)
        start = torch.cuda.Event(enable_timing=True)
        stop = torch.cuda.Event(enable_timing=True)
        start.record()
        rc = self._call(name, *args)
        stop.record()
        self.timers.setdefault(key, []).append((start, stop))
        return rc

    def timer_summary(self):
        """{key: (launch groups, mean ms)} -- call after a synchronize."""
        out = {}
        for key, pairs in (self.timers or {}).items():
            ms = [a.elapsed_time(b) for a, b in pairs]
            out[key] = (len(ms), sum(ms) / max(len(ms), 1))
        return out

    def host_tick(self, label, t0):
        """Accumulates host wall time per label (diagnostics of the fit's host path)."""
        t1 = time.perf_counter()
        self.host_times[label] = self.host_times.get(label, 0.0) + t1 - t0
        # thread CPU time next to it tells "computing" from "blocked" (GIL, driver lock, sync)
        c1 = time.thread_time()
        c0 = self._cpu_mark.get(threading.get_ident(), c1)
        self._cpu_mark[threading.get_ident()] = c1
        self.host_times[label + ":cpu"] = self.host_times.get(label + ":cpu", 0.0) + c1 - c0
        return t1

    def launch_count(self):
        return _lib.launch_count(self._lib)

    def stream(self):
        if self.device.type == "cuda":
            return torch.cuda.current_stream(self.device).cuda_stream
        return None

    def tensor(self, array, dtype):
        # always a fresh buffer: in-place kernels must never alias the caller's host array
        t = torch.as_tensor(np.ascontiguousarray(array))
        return t.to(device=self.device, dtype=dtype, copy=True).contiguous()

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)

    def zeros(self, shape, dtype):
        return torch.zeros(shape, dtype=dtype, device=self.device)

    def _check(self, t, dtype, name):
        if t.dtype != dtype or not t.is_contiguous() or t.device != self.device:
            raise ValueError(f"{name}: need contiguous {dtype} tensor on {self.device}, got "
                             f"{t.dtype} contiguous={t.is_contiguous()} on {t.device}")

    # ------------------------------------------------------------------ KL setup
    @staticmethod
    def full_mask(ndir):
        m = (1 << ndir) - 1
        lo, hi = m & (2**64 - 1), m >> 64
        # torch has no uint64 arithmetic: masks travel as int64 bit patterns
        return [lo - 2**64 if lo >= 2**63 else lo, hi - 2**64 if hi >= 2**63 else hi]

    def kl_decompose(self, pp, masks, r0=R_0, beta=BETA, cut=PINV_CUT, out=None):
        """ssf_kl_decompose: pp [ndir, 2] f64, masks [npat, 2] int64 bit patterns.
        Returns C, pinvC, U [npat, ndir, ndir] and sv [npat, ndir] (``out``: the same four as
        preallocated contiguous tensors)."""
        self._check(pp, torch.float64, "pp")
        self._check(masks, torch.int64, "masks")
        ndir, npat = pp.shape[0], masks.shape[0]
        if out is not None:
            c, p, u, sv = out
            for t, nm in ((c, "C"), (p, "pinvC"), (u, "U"), (sv, "sv")):
                self._check(t, torch.float64, nm)
                if t.shape[0] != npat:
                    raise ValueError("decompose output has the wrong number of patterns")
        else:
            c = self.empty((npat, ndir, ndir), torch.float64)
            p = self.empty((npat, ndir, ndir), torch.float64)
            u = self.empty((npat, ndir, ndir), torch.float64)
            sv = self.empty((npat, ndir), torch.float64)
        self._timed_call("kl_decompose", "ssf_kl_decompose", _ptr(pp), ndir, _ptr(masks), npat, r0, beta, cut,
                  _ptr(c), _ptr(p), _ptr(u), _ptr(sv), self.stream())
        return c, p, u, sv

    def kl_flag_outliers(self, resid, weights, nsigma):
        """ssf_kl_flag_outliers in place on ``weights``; returns the [nsys, 2] masks.
        ``nsigma = inf`` flags nothing and just reports the masks of the weights."""
        nsys, ndir = resid.shape
        self._check(resid, torch.float64, "resid")
        self._check(weights, torch.float32, "weights")
        masks = self.empty((nsys, 2), torch.int64)
        self._timed_call("kl_flag_outliers", "ssf_kl_flag_outliers", _ptr(resid), _ptr(weights), ndir, nsys, float(nsigma),
                  _ptr(masks), self.stream())
        return masks

    def kl_flag_outliers_amp(self, val, resid, weights, ngroups, nsigma):
        """ssf_kl_flag_outliers_amp in place on ``weights`` (amplitude type: one rms per
        (freq, station, pol) group over all directions and times); returns the masks.
        ``val=None``: tec type, the residual itself is clipped instead of the log residual."""
        nsys, ndir = resid.shape
        if val is not None:
            self._check(val, torch.float64, "val")
        self._check(resid, torch.float64, "resid")
        self._check(weights, torch.float32, "weights")
        masks = self.empty((nsys, 2), torch.int64)
        std = self.empty((ngroups,), torch.float64)
        self._timed_call("kl_flag_outliers", "ssf_kl_flag_outliers_amp", _ptr(val), _ptr(resid), _ptr(weights),
                         ndir, nsys, ngroups, float(nsigma), _ptr(std), _ptr(masks), self.stream())
        return masks

    def kl_fit_pass(self, phase, weights, pat, pp, mats, mode, order, station_order, white, resid,
                    r0=R_0, beta=BETA, cut=PINV_CUT, status=None):
        nsys, ndir = phase.shape
        c, p, u = mats
        for t, dt, nm in ((phase, torch.float64, "phase"), (weights, torch.float32, "weights"),
                          (pat, torch.int32, "pat"), (order, torch.float64, "order"),
                          (station_order, torch.float64, "station_order"),
                          (white, torch.float64, "white"), (resid, torch.float64, "resid")):
            self._check(t, dt, nm)
        if status is None:
            status = self.zeros((1,), torch.int32)
        self._timed_call("kl_fit_pass", "ssf_kl_fit_pass", _ptr(phase), _ptr(weights), _ptr(pat), _ptr(pp), ndir, nsys,
                   _ptr(c), _ptr(p), _ptr(u), r0, beta, cut, mode, _ptr(order), _ptr(station_order),
                   _ptr(white), _ptr(resid), _ptr(status), self.stream())
        return status

    def kl_fit(self, phase, weights, station_order, pp, niter=2, nsigma=5.0, adjust_order=True,
               r0=R_0, beta=BETA, cut=PINV_CUT, cache=None, check=True, screen_type="phase",
               ngroups=None):
        """The iteration structure of _process_station (stationscreen.py:658-779) for a batch
        of independent systems.  phase [nsys, ndir] f64 (already referenced), weights
        [nsys, ndir] f32 (modified in place by the flagging), station_order [nsys] f64.
        Returns (white, resid, order) with shapes [nsys, ndir], [nsys, ndir], [nsys].
        ``cache``: a PatternCache for these pierce points (decompositions of flag patterns are
        kept across calls); a temporary one is made if absent.
        ``check=False`` defers the convergence / weight-range checks: the caller must call
        ``check_fit_status()`` later.
        ``screen_type="amplitude"``: ``phase`` holds amplitudes, rows are time-major with
        ``ngroups`` = freq * station * pol systems per time slot (the outlier rms is taken per
        group over all times, stationscreen.py:341-346); ``white`` are log10 screens.
        ``screen_type="tec"``: the values themselves are fitted (:549-562), clipped like the
        amplitude type but on the plain residual (:338-346, :661-663), plain chi2 (:742-746)."""
        nsys, ndir = phase.shape
        amp = screen_type == "amplitude"
        tec = screen_type == "tec"
        if screen_type not in ("phase", "amplitude", "tec"):
            raise ValueError("screen_type must be 'phase', 'amplitude' or 'tec'")
        if (amp or tec) and (not ngroups or nsys % ngroups):
            raise ValueError("amplitude / tec fits need ngroups (systems per time slot) dividing nsys")
        t0 = time.perf_counter()
        white = self.zeros((nsys, ndir), torch.float64)
        resid = self.zeros((nsys, ndir), torch.float64)
        order = station_order.clone()
        self.host_tick("fit_alloc", t0)
        if cache is None:
            cache = PatternCache(self, pp, r0, beta, cut, max_bytes=256 << 20)
        self._caches.add(cache)
        if self._fit_status is None:
            self._fit_status = self.zeros((2,), torch.int32)     # [weights out of range, no convergence]
        for it in range(niter):
            t0 = time.perf_counter()
            if amp or tec:
                masks = self.kl_flag_outliers_amp(phase if amp else None, resid, weights, ngroups,
                                                  nsigma if it > 0 else math.inf)
            else:
                masks = self.kl_flag_outliers(resid, weights, nsigma if it > 0 else math.inf)
            t0 = self.host_tick("flag_launch", t0)
            if it > 0 and not adjust_order:
                # the flagged weights alias the "previous" weights in the reference, so
                # without order adjustment nothing is refitted (stationscreen.py:694-701)
                continue
            mode = (1 if (it > 0 and adjust_order) else 0) | (2 if amp else 0) | (4 if tec else 0)
            # one lookup + fit pass per run of at most max_slots - 1 systems: even if every one
            # of them had a flag pattern of its own the cache cannot run out of slots inside a
            # pass (between passes it forgets all patterns but the full one when it has to)
            step = cache.max_slots - 1
            for s0 in range(0, nsys, step):
                s1 = min(nsys, s0 + step)
                pat = cache.lookup(masks[s0:s1])
                t0 = self.host_tick("pattern_lookup", t0)
                self.kl_fit_pass(phase[s0:s1], weights[s0:s1], pat, pp, cache.mats, mode, order[s0:s1],
                                 station_order[s0:s1], white[s0:s1], resid[s0:s1], r0, beta, cut,
                                 status=self._fit_status[0:1])
                t0 = self.host_tick("fit_pass_launch", t0)
        if check:
            self.check_fit_status()
        return white, resid, order

    def check_fit_status(self):
        if self._fit_status is None:
            return
        bad_w, no_conv = (int(v) for v in self._fit_status.tolist())
        self._fit_status.zero_()
        for cache in list(self._caches):
            if cache.counters[2].item():
                cache.counters[2:3].zero_()
                raise _lib.SsfError("ssf_kl_patterns_update", -5, "one lookup held more distinct flag patterns "
                                    "than the pattern cache has slots (systems without a slot were not "
                                    "fitted): pass at most max_slots - 1 systems per lookup")
        if no_conv:
            raise _lib.SsfError("ssf_kl_decompose", -3, "Jacobi eigen-solver did not converge")
        if bad_w:
            raise _lib.SsfError("ssf_kl_fit_pass", -4, "a system has a weight in (0, pinv_cut]: "
                                "outside the normal-equation path of this kernel")

    # ------------------------------------------------------------------ KL evaluation
    def kl_basis(self, x_coord, y_coord, pp, pinv_c, r0=R_0, beta=BETA):
        nx, ny, ndir = x_coord.numel(), y_coord.numel(), pp.shape[0]
        bt = self.empty((ndir, ny * nx), torch.float32)
        self._call("ssf_kl_basis", _ptr(x_coord), _ptr(y_coord), nx, ny, _ptr(pp), ndir,
                  _ptr(pinv_c), r0, beta, _ptr(bt), self.stream())
        return bt

    def kl_coeffs(self, white, c, out=None):
        ncol, ndir = white.shape
        self._check(white, torch.float64, "white")
        ldg = (ndir + 3) // 4 * 4
        g = out if out is not None else self.empty((ncol, ldg), torch.float32)
        self._call("ssf_kl_coeffs", _ptr(white), ncol, ndir, _ptr(c), _ptr(g), ldg, self.stream())
        return g

    @staticmethod
    def out_planes(out_flags):
        """Planes per column a cube kernel stores for these SSF_OUT_* flags."""
        return 2 if out_flags & _lib.OUT_PLANES2 else 4

    def kl_eval(self, bt, g, ndir, slab, variant=-1, out_flags=0):
        """slab: float32 [ncol, 4, npix], or [ncol, 2, npix] with OUT_PLANES2."""
        ncol, ldg = g.shape
        npix = bt.shape[1]
        self._check(slab, torch.float32, "slab")
        if slab.numel() != ncol * self.out_planes(out_flags) * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("kl_eval", "ssf_kl_eval", _ptr(bt), _ptr(g), ndir, ldg, npix, ncol, _ptr(slab), variant,
                         out_flags, self.stream())
        return slab

    def kl_eval_tec(self, bt, g, ndir, slab, variant=-1, out_flags=0):
        """dTEC screens: slab float32 [ncol, npix] = the screen values themselves."""
        ncol, ldg = g.shape
        npix = bt.shape[1]
        self._check(slab, torch.float32, "slab")
        if slab.numel() != ncol * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("kl_eval", "ssf_kl_eval_tec", _ptr(bt), _ptr(g), ndir, ldg, npix, ncol, _ptr(slab), variant,
                         out_flags, self.stream())
        return slab

    def kl_eval_gain(self, bt, g3, ndir, slab, variant=-1, out_flags=0):
        """g3 [ncol, 3, ldg] f32 (phase, Axx, Ayy coefficients) -> gain cube slab."""
        ncol, _, ldg = g3.shape
        npix = bt.shape[1]
        self._check(slab, torch.float32, "slab")
        self._check(g3, torch.float32, "g3")
        if slab.numel() != ncol * 4 * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("kl_eval", "ssf_kl_eval_gain", _ptr(bt), _ptr(g3), ndir, ldg, npix, ncol, _ptr(slab),
                         variant, out_flags, self.stream())
        return slab

    # ------------------------------------------------------------------ Voronoi
    def voronoi_assign(self, px, py, nx, ny):
        tmpl = self.empty((ny, nx), torch.int32)
        self._call("ssf_voronoi_assign", _ptr(px), _ptr(py), px.numel(), nx, ny, _ptr(tmpl),
                  self.stream())
        return tmpl

    def voronoi_lut(self, phase, amp=None, out=None):
        ncol, ndir = phase.shape
        self._check(phase, torch.float64, "phase")
        lut = out if out is not None else self.empty((ncol, ndir, 2), torch.float32)
        self._call("ssf_voronoi_lut", _ptr(phase), _ptr(amp), ncol, ndir, _ptr(lut), self.stream())
        return lut

    def voronoi_fill(self, tmpl, lut, slab, fix_nan=True, out_flags=0):
        ncol, ndir, _ = lut.shape
        npix = tmpl.numel()
        self._check(tmpl, torch.int32, "tmpl")
        self._check(slab, torch.float32, "slab")
        if slab.numel() != ncol * self.out_planes(out_flags) * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("voronoi_fill", "ssf_voronoi_fill", _ptr(tmpl), _ptr(lut), ndir, npix, ncol, _ptr(slab),
                         1 if fix_nan else 0, out_flags, self.stream())
        return slab

    def voronoi_lut_gain(self, phase, amp):
        """phase [ncol, ndir] f64, amp [ncol, ndir, 2] f64 (XX, YY) -> lut4 [ncol, ndir, 4] f32."""
        ncol, ndir = phase.shape
        self._check(phase, torch.float64, "phase")
        self._check(amp, torch.float64, "amp")
        if tuple(amp.shape) != (ncol, ndir, 2):
            raise ValueError("amp must be [ncol, ndir, 2]")
        lut = self.empty((ncol, ndir, 4), torch.float32)
        self._call("ssf_voronoi_lut_gain", _ptr(phase), _ptr(amp), ncol, ndir, _ptr(lut), self.stream())
        return lut

    def voronoi_fill_gain(self, tmpl, lut4, slab, fix_nan=True, out_flags=0):
        ncol, ndir, _ = lut4.shape
        npix = tmpl.numel()
        self._check(tmpl, torch.int32, "tmpl")
        self._check(slab, torch.float32, "slab")
        if slab.numel() != ncol * 4 * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("voronoi_fill", "ssf_voronoi_fill_gain", _ptr(tmpl), _ptr(lut4), ndir, npix, ncol,
                         _ptr(slab), 1 if fix_nan else 0, out_flags, self.stream())
        return slab

    def voronoi_indicators(self, tmpl, ndir):
        npix = tmpl.numel()
        ind = self.empty((ndir,) + tuple(tmpl.shape), torch.float32)
        self._call("ssf_voronoi_indicators", _ptr(tmpl), ndir, npix, _ptr(ind), self.stream())
        return ind

    def voronoi_contract(self, smoothed_ind, lut, slab, variant=-1, out_flags=0):
        ncol, ndir, _ = lut.shape
        npix = smoothed_ind.numel() // ndir
        if slab.numel() != ncol * self.out_planes(out_flags) * npix:
            raise ValueError("slab has the wrong size")
        self._timed_call("voronoi_contract", "ssf_voronoi_contract", _ptr(smoothed_ind), _ptr(lut), ndir, npix, ncol,
                         _ptr(slab), variant, out_flags, self.stream())
        return slab

    def gauss_smooth(self, images, sigma, scratch=None):
        """In place on images [..., ny, nx] float32."""
        self._check(images, torch.float32, "images")
        ny, nx = images.shape[-2:]
        nimg = images.numel() // (ny * nx)
        if scratch is None:
            scratch = torch.empty_like(images)
        self._timed_call("gauss_smooth", "ssf_gauss_smooth", _ptr(images), nimg, ny, nx, float(sigma), _ptr(scratch),
                  self.stream())
        return images

    def gauss_smooth_static(self, images, sigma):
        """gauss_smooth with the library's own fixed scratch (``ssf_gauss_smooth_static``): no
        allocation anywhere in the call, at most ``ssf_gauss_smooth_static_max_floats()`` pixels,
        one call at a time per GPU (the scratch is shared).  Not used by the screens."""
        self._check(images, torch.float32, "images")
        ny, nx = images.shape[-2:]
        nimg = images.numel() // (ny * nx)
        self._call("ssf_gauss_smooth_static", _ptr(images), nimg, ny, nx, float(sigma), self.stream())
        return images

    def gauss_smooth_slab(self, slab, sigma, fix_nan=True, scratch=None):
        """In place on a slab [ncol, 4, ny, nx] float32; ``fix_nan`` fuses the NaN fix of
        screen.py:364-378 into the second pass."""
        self._check(slab, torch.float32, "slab")
        ncol, planes, ny, nx = slab.shape
        if planes != 4:
            raise ValueError("slab must be [ncol, 4, ny, nx]")
        if scratch is None:
            scratch = torch.empty_like(slab)
        self._timed_call("gauss_smooth", "ssf_gauss_smooth_slab", _ptr(slab), ncol, ny, nx, float(sigma),
                         _ptr(scratch), 1 if fix_nan else 0, self.stream())
        return slab

    def nan_fix(self, slab, ncol, npix):
        self._call("ssf_nan_fix", _ptr(slab), ncol, npix, self.stream())
        return slab


    def bswap32(self, t):
        """In place: float32 little-endian -> big-endian (FITS storage order)."""
        self._call("ssf_bswap32", _ptr(t), t.numel(), self.stream())
        return t

    def copy_rows_d2h(self, dst_host, dst_pitch, src, src_pitch, width_bytes, rows):
        """ssf_copy_rows_d2h on the current stream: ``rows`` runs of ``width_bytes`` from the
        device tensor ``src`` to the (pinned) host tensor / address ``dst_host``."""
        dst = dst_host if isinstance(dst_host, int) else dst_host.data_ptr()
        self._call("ssf_copy_rows_d2h", dst, dst_pitch, _ptr(src), src_pitch, width_bytes, rows, self.stream())


class PatternCache:
    """Decompositions (C_S, pinv C_S, U_S) of the flag patterns seen so far for one set of
    pierce points, kept on the device across fit calls -- together with the map from a
    pattern to its slot (``ssf_kl_patterns_update``: device hash table, atomic slot counter,
    new slots decomposed in the same launch sequence).

    The reference recomputes the SVD for every system with a flagged direction
    (stationscreen.py:493-499); systems that share a flag pattern share the result, within a
    slab and across slabs.  Nothing here reads anything back: a fit pass is a pure sequence of
    stream-ordered launches, so the host never waits for the GPU and the fit of the next slabs
    can be issued while the evaluation kernel owns the device.  Slot 0 is the full pattern.
    When the slots run out the device forgets all patterns but slot 0 and starts over."""

    def __init__(self, engine, pp, r0=R_0, beta=BETA, cut=PINV_CUT, max_bytes=None):
        if max_bytes is None:
            # a sixteenth of the free device memory, 2 .. 16 GiB: when the cache is full it forgets
            # everything (and the next passes decompose every pattern again -- 8 ms per 2048
            # matrices of 100 directions), so on a 180 GB part it should rarely be full: at the
            # SKA-Low shape (one new pattern per ~4 systems at 1 % flags) 16 GiB last ~500 slabs
            free = torch.cuda.mem_get_info(engine.device)[0] if engine.device.type == "cuda" else 0
            max_bytes = int(min(max(free // 16, 2 << 30), 16 << 30)) if free else 2 << 30
        self.eng = engine
        self.pp = pp
        self.ndir = n = pp.shape[0]
        self.params = (r0, beta, cut)
        per = (3 * n * n + n) * 8 + 16
        self.max_slots = int(max(2, min(max_bytes // per, 1 << 16)))
        self.hash_cap = 1 << int(math.ceil(math.log2(2 * self.max_slots)))
        self.c, self.p, self.u = (engine.empty((self.max_slots, n, n), torch.float64) for _ in range(3))
        self.sv = engine.empty((self.max_slots, n), torch.float64)
        self.slot_masks = engine.zeros((self.max_slots, 2), torch.int64)
        self.table_keys = engine.zeros((self.hash_cap, 2), torch.int64)
        self.table_slots = torch.full((self.hash_cap,), -1, dtype=torch.int32, device=engine.device)
        self.counters = torch.tensor([1, 1, 0, 0], dtype=torch.int32, device=engine.device)
        full = torch.tensor([Engine.full_mask(n)], dtype=torch.int64, device=engine.device)
        self.slot_masks[0:1].copy_(full)
        engine.kl_decompose(pp, full, r0, beta, cut,
                            out=(self.c[0:1], self.p[0:1], self.u[0:1], self.sv[0:1]))
        if engine._fit_status is None:
            engine._fit_status = engine.zeros((2,), torch.int32)
        engine._fit_status[1:2] += torch.isnan(self.sv[0]).any().to(torch.int32)
        engine._caches.add(self)                      # check_fit_status reports this cache's overflow flag

    @property
    def mats(self):
        return self.c, self.p, self.u

    def lookup(self, masks, out=None):
        """masks: device [nsys, 2] int64 bit patterns -> device int32 [nsys] slot of each
        system's pattern; the patterns not seen before are decomposed (stream-ordered)."""
        eng = self.eng
        nsys = masks.shape[0]
        pat = out if out is not None else eng.empty((nsys,), torch.int32)
        r0, beta, cut = self.params
        eng._timed_call("kl_patterns_update", "ssf_kl_patterns_update", _ptr(masks), nsys, self.ndir,
                        _ptr(self.pp), r0, beta, cut, _ptr(self.table_keys), _ptr(self.table_slots),
                        self.hash_cap, _ptr(self.slot_masks), self.max_slots, _ptr(self.c), _ptr(self.p),
                        _ptr(self.u), _ptr(self.sv), _ptr(self.counters), _ptr(eng._fit_status[1:2]),
                        _ptr(pat), eng.stream())
        return pat

    def stats(self):
        """(slots in use, overflow flag) -- reads the counters back (synchronises)."""
        c = self.counters.tolist()
        return c[0], c[2]

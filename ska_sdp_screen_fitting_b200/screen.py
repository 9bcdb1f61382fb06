"""
Base class of the a-term screen generators -- the drop-in for the reference's
``screen.py:19-411`` (same constructor, attributes and methods), with the per-(freq,
station) Python write loop replaced by GPU time slabs:

    reference                                     here
    ---------                                     ----
    for freq: for station:                        for each time slab [t0, t1):
        make_matrix(...)  (numpy / Pool)              one device slab [t, f, a, 4, y, x]
        gaussian_filter per time                      produced by CUDA kernels (subclass),
    NaN fix over the cube                             smoothing + NaN fix on the device,
    hdu.writeto (second full write)                   byte-swapped on the device and written
                                                      once at its offset in the FITS file

With ``torch.distributed`` initialised, every rank owns a contiguous range of the time
slabs of each output file and writes its own byte range; only file-level metadata is
exchanged.
"""
import logging
import os

import numpy as np
import torch

from . import fits, wcs
from .engine import Engine


def _dist_info():
    try:
        import torch.distributed as dist
    except ImportError:  # pragma: no cover
        return 0, 1, None
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size(), dist
    return 0, 1, None


def split_range(start, stop, parts, index):
    """Contiguous near-equal partition of [start, stop) into ``parts``; returns part ``index``."""
    n = stop - start
    base, extra = divmod(n, parts)
    lo = start + index * base + min(index, extra)
    return lo, lo + base + (1 if index < extra else 0)


def _interp_axis(x_old, values, axis, x_new, kind):
    """``scipy.interpolate.interp1d(x_old, values, axis=axis, kind=kind,
    fill_value="extrapolate")(x_new)`` (screen.py:134-150).  "nearest" is restated (sorted
    abscissae, ties at a midpoint go to the lower sample, ends clamp -- scipy's rule); any other
    kind is handed to scipy itself."""
    x_old = np.asarray(x_old, np.float64)
    x_new = np.asarray(x_new, np.float64)
    if kind == "nearest":
        order = np.argsort(x_old, kind="mergesort")
        xs = x_old[order]
        mids = 0.5 * (xs[1:] + xs[:-1])
        idx = np.clip(np.searchsorted(mids, x_new, side="left"), 0, len(xs) - 1)
        return np.take(np.take(values, order, axis=axis), idx, axis=axis)
    from scipy import interpolate as si

    return si.interp1d(x_old, values, axis=axis, kind=kind, fill_value="extrapolate")(x_new)


def time_chunks(times):
    """Output-file boundaries (screen.py:283-294): a new file wherever the time step exceeds
    1.2 x the smallest step.  Returns the list of stop indices.  The reference's further,
    free-memory dependent splits (screen.py:296-317) are deliberately not reproduced: slabs
    are streamed, so memory never forces a file boundary and the file list is deterministic."""
    times = np.asarray(times)
    if len(times) > 2:
        delta = times[1:] - times[:-1]
        gaps = np.where(delta > np.min(delta) * 1.2)[0] + 1
        return [int(g) for g in gaps] + [len(times)]
    return [len(times)]


class Screen:
    """Master class for a-term screens (reference: screen.py:19-99).

    Parameters are those of the reference: ``name`` (output root), the h5parm and sky model
    file names, image centre ``rad``/``dec`` (deg), ``width_ra``/``width_dec`` (deg; the image
    is forced square), solset / soltab names."""

    def __init__(self, name, h5parm_filename, skymodel_filename, rad, dec, width_ra, width_dec,
                 solset_name="sol000", phase_soltab_name="phase000", amplitude_soltab_name=None):
        self.name = name
        self.log = logging.getLogger(f"rapthor:{self.name}")
        self.input_h5parm_filename = h5parm_filename
        self.input_skymodel_filename = skymodel_filename
        self.input_solset_name = solset_name
        self.input_phase_soltab_name = phase_soltab_name
        self.input_amplitude_soltab_name = amplitude_soltab_name
        self.phase_only = amplitude_soltab_name is None
        if isinstance(rad, str) or isinstance(dec, str):
            from .skymodel import dec_to_deg, ra_to_deg
            rad = ra_to_deg(rad) if isinstance(rad, str) else rad
            dec = dec_to_deg(dec) if isinstance(dec, str) else dec
        self.rad = rad
        self.dec = dec
        width = max(width_ra, width_dec)            # square images only (screen.py:76-80)
        self.width_ra = width
        self.width_dec = width
        self.log_amps = False
        self.times_amp = None
        self.times_ph = []
        self.vals_amp = None
        self.vals_ph = None
        self.freqs_amp = None
        self.freqs_ph = None
        self.station_names = None
        self.source_names = None
        self.source_dict = None
        self.source_positions = None
        self.station_dict = None
        self.station_positions = None
        self.ncpu = None
        # --- engine-side state (not part of the reference's attribute set)
        self.solutions = None            # optional in-memory h5parm_io.Solutions (bypasses file)
        self.slab_bytes = 2 << 30        # device slab budget for write()
        self.timings = {}
        self._engine = None
        self._dev_vals = None            # device [t, f, a, d] float64: what make_matrix consumes
        self._dev_amp = None             # device [t, f, a, d, 2] float64 (gain solutions only)

    # ------------------------------------------------------------------ plumbing
    @property
    def engine(self):
        if self._engine is None:
            self._engine = Engine()
        return self._engine

    def _load_solutions(self, soltab_name):
        if isinstance(self.solutions, dict):         # {soltab name: Solutions} (phase + amplitude)
            return self.solutions[soltab_name]
        if self.solutions is not None:
            return self.solutions
        from .h5parm_io import read_solutions
        return read_solutions(self.input_h5parm_filename, self.input_solset_name, soltab_name)

    def image_shape(self, cellsize_deg):
        n_x = wcs.image_size(self.width_ra, cellsize_deg)
        n_y = wcs.image_size(self.width_dec, cellsize_deg)
        return n_y, n_x

    # ------------------------------------------------------------------ reference API
    def fit(self):
        """Fits screens to the input solutions (implemented in the subclasses)."""

    def interpolate(self, interp_kind="nearest"):
        """screen.py:108-154: resample the slow amplitudes (in log space) onto the fast-phase
        time and frequency grid; phase-only solutions return immediately (:118-119).  Host
        marshalling, once per run; the result is uploaded as the device tensor the cube
        kernels read ([time, freq, ant, dir, pol])."""
        if self.phase_only:
            return
        vals_amp = np.asarray(self.vals_amp)
        n_t, n_f = self.vals_ph.shape[0], self.vals_ph.shape[1]
        if len(self.times_amp) == 1:
            # a single time: the values are repeated as needed (numpy.resize semantics, :121-126)
            new_shape = list(vals_amp.shape)
            new_shape[0], new_shape[1] = n_t, n_f
            vals_amp = np.resize(vals_amp, new_shape)
        else:
            logvals = vals_amp if self.log_amps else np.log10(vals_amp)
            if vals_amp.shape[0] != n_t:
                logvals = _interp_axis(self.times_amp, logvals, 0, self.times_ph, interp_kind)
            if vals_amp.shape[1] != n_f:
                logvals = _interp_axis(self.freqs_amp, logvals, 1, self.freqs_ph, interp_kind)
            vals_amp = logvals if self.log_amps else 10 ** logvals
        self.vals_amp = vals_amp
        self._dev_amp = self.engine.tensor(np.ascontiguousarray(vals_amp[..., :2]), torch.float64)

    def get_memory_usage(self, cellsize_deg):
        """GB of HOST memory per time slot.  The reference's estimates (kl_screen.py:157-190,
        voronoi_screen.py:104-130) budget for whole-cube residency; here a slot only ever
        exists as part of a streamed slab."""
        n_y, n_x = self.image_shape(cellsize_deg)
        return len(self.freqs_ph) * len(self.station_names) * 4 * n_y * n_x * 4 / 1024 ** 3

    def make_matrix(self, t_start_index, t_stop_index, freq_ind, stat_ind, cellsize_deg, out_dir, ncpu):
        """Returns float64 [t_stop - t_start, 4, ny, nx] for one (freq, station), computed by
        the same device path as write() (reference: screen.py:197-245)."""
        n_y, n_x = self.image_shape(cellsize_deg)
        cols = self._dev_vals[t_start_index:t_stop_index, freq_ind, stat_ind, :].contiguous()
        amp = None
        if not self.phase_only:
            amp = self._dev_amp[t_start_index:t_stop_index, freq_ind, stat_ind].contiguous()
        slab = self.engine.empty((cols.shape[0], 4, n_y, n_x), torch.float32)
        self._fill_columns(cols, slab, cellsize_deg, out_dir, smooth_pix=0.0, fix_nan=False, amp=amp)
        return slab.double().cpu().numpy()

    def _fill_columns(self, cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan, amp=None):
        """cols: device [ncol, ndir] float64; amp: device [ncol, ndir, 2] float64 (XX, YY) for
        gain solutions; slab: device float32 [ncol, 4, ny, nx]."""
        raise NotImplementedError

    def _prepare(self, cellsize_deg, out_dir):
        """Once-per-run setup before slabs are produced (basis / template)."""

    def process(self, ncpu=0):
        """Fit, then interpolate (screen.py:396-411)."""
        self.ncpu = ncpu
        self.fit()
        self.interpolate()

    # ------------------------------------------------------------------ write
    def slab_columns(self, cellsize_deg):
        """(time, freq, antenna) columns per device slab.  A slab is a run of whole columns, not
        of whole time slots: one SKA-Low time slot (32 freqs x 512 stations x 2048^2) is 1.1 TB."""
        n_y, n_x = self.image_shape(cellsize_deg)
        return max(1, int(self.slab_bytes // (16 * n_y * n_x)))

    def write(self, out_dir, cellsize_deg, smooth_pix=0, ncpu=0):
        """Writes ``{name}_{g}.fits`` per gap-free time chunk and ``{name}.txt``
        (screen.py:260-394)."""
        self.ncpu = ncpu
        rank, world, dist = _dist_info()
        eng = self.engine
        n_y, n_x = self.image_shape(cellsize_deg)
        n_f, n_a = len(self.freqs_ph), len(self.station_names)
        times = np.asarray(self.times_ph, np.float64)
        self._prepare(cellsize_deg, out_dir)
        step = self.slab_columns(cellsize_deg)
        per_t = n_f * n_a
        n_d = self._dev_vals.shape[-1]
        flat_vals = self._dev_vals.reshape(-1, n_d)
        flat_amp = None if self.phase_only else self._dev_amp.reshape(-1, n_d, 2)
        pinned = [None, None]
        outfiles = []
        g_start = 0
        voxels = 0
        for gnum, g_stop in enumerate(time_chunks(times)):
            outfile = os.path.join(out_dir, f"{self.name}_{gnum}.fits")
            shape = (g_stop - g_start, n_f, n_a, 4, n_y, n_x)
            cards = fits.aterm_cards(shape, self.rad, self.dec, cellsize_deg, self.freqs_ph,
                                     times[g_start:g_stop])
            if rank == 0:
                writer = fits.CubeWriter(outfile, shape, cards)
            if dist is not None:
                dist.barrier()
            if rank != 0:
                writer = fits.CubeWriter(outfile, shape, cards, create=False)
            my0, my1 = split_range(g_start, g_stop, world, rank)
            pending = None
            buf = 0
            for c0 in range(my0 * per_t, my1 * per_t, step):
                c1 = min(my1 * per_t, c0 + step)
                cols = flat_vals[c0:c1]
                amp = None if flat_amp is None else flat_amp[c0:c1]
                slab = eng.empty((cols.shape[0], 4, n_y, n_x), torch.float32)
                self._fill_columns(cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan=True, amp=amp)
                eng.bswap32(slab)
                if pinned[buf] is None or pinned[buf].numel() < slab.numel():
                    pinned[buf] = torch.empty(slab.numel(), dtype=torch.float32,
                                              pin_memory=eng.device.type == "cuda")
                host = pinned[buf][:slab.numel()]
                host.copy_(slab.view(-1), non_blocking=True)
                event = None
                if eng.device.type == "cuda":
                    event = torch.cuda.Event()
                    event.record()
                if pending is not None:
                    self._flush(writer, *pending)
                pending = (c0 - g_start * per_t, host, event)
                buf ^= 1
                voxels += cols.shape[0] * n_y * n_x
            if pending is not None:
                self._flush(writer, *pending)
            writer.close()
            outfiles.append(outfile)
            g_start = g_stop
        if dist is not None:
            dist.barrier()
        self.timings["voxels_written"] = voxels
        if rank == 0:
            with open(os.path.join(out_dir, f"{self.name}.txt"), "w", encoding="utf8") as handle:
                handle.writelines([o + "\n" for o in outfiles])
        return outfiles

    @staticmethod
    def _flush(writer, col_rel, host, event):
        if event is not None:
            event.synchronize()
        writer.write_columns(col_rel, host.numpy().view(">f4"), already_big_endian=True)

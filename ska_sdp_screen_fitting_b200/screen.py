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
import queue
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

from . import _lib, fits, wcs
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


def time_chunks(times, max_times_per_file=None):
    """Output-file boundaries (screen.py:283-294): a new file wherever the time step exceeds
    1.2 x the smallest step.  Returns the list of stop indices.

    The reference then halves any chunk longer than ``available memory / get_memory_usage``
    time slots, repeatedly (screen.py:296-317): its file list depends on the free memory of the
    machine it ran on.  Slabs are streamed here, so memory never forces a boundary and by
    default the file list is a function of the time axis alone; ``max_times_per_file`` (the
    reference's ``max_ntimes``) reproduces its extra splits for a given budget, with the same
    first-oversized-chunk-first halving order."""
    times = np.asarray(times)
    if len(times) > 2:
        delta = times[1:] - times[:-1]
        gaps = np.where(delta > np.min(delta) * 1.2)[0] + 1
        stops = [int(g) for g in gaps] + [len(times)]
        if max_times_per_file:
            limit = max(1, int(max_times_per_file))
            again = True
            while again:
                again = False
                start = 0
                for k, stop in enumerate(list(stops)):
                    if stop - start > limit:
                        stops.insert(k, start + (stop - start) // 2)
                        again = True
                        break
                    start = stop
        return stops
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
        self.slab_bytes = 1 << 30        # device slab budget for write() (two slabs are in flight)
        self.block_columns = 24576       # columns handed to write() at a time by _blocks
        self.host_buffers = 3            # pinned staging buffers between the copy stream and the file
        self.write_threads = min(8, os.cpu_count() or 1)    # placement threads per staged slab (tmpfs page allocation
                                                            # stops scaling at 8: profiles/r02_file_sink.txt)
        self.max_times_per_file = None   # reference's memory-driven extra file splits (screen.py:296-317), opt-in
        self.preallocate = False         # fallocate thread ahead of the placement (no gain on tmpfs: profiles/r02_file_sink.txt)
        self.file_mapping = True         # place slabs by threaded copies into a mapping of the file (else pwritev)
        self.sink = "file"               # "null": stop at the pinned host buffer (benchmarks: link vs file system)
        self.aterm_type = "gain"         # "tec": 5-axis dTEC cube (KLScreen on a soltab of type tec)
        self.timings = {}
        self._chunks = None              # [(g_start, g_stop, my0, my1, local0)] from _plan_rows
        self._copy_stream = None
        self._rings = None
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
        n_t, n_f = len(self.times_ph), len(self.freqs_ph)
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
        # the cube kernels read this rank's time rows only
        local = vals_amp if self._chunks is None or _dist_info()[1] == 1 else vals_amp[self._local_rows()]
        self._dev_amp = self.engine.tensor(np.ascontiguousarray(local[..., :2]), torch.float64)

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

    def _fill_columns(self, cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan, amp=None, out_flags=0):
        """cols: device [ncol, ndir] float64; amp: device [ncol, ndir, 2] float64 (XX, YY) for
        gain solutions; slab: device float32 [ncol, 4, ny, nx] ([ncol, 2, ny, nx] with
        OUT_PLANES2 in ``out_flags``, one of the formats ``_out_format`` offered)."""
        raise NotImplementedError

    def _prepare(self, cellsize_deg, out_dir):
        """Once-per-run setup before slabs are produced (basis / template)."""

    def process(self, ncpu=0):
        """Fit, then interpolate (screen.py:396-411)."""
        self.ncpu = ncpu
        self.fit()
        self.interpolate()

    # ------------------------------------------------------------------ time rows of this rank
    def _plan_rows(self, times):
        """Output-file chunks (screen.py:283-294) and this rank's share of each: with
        ``torch.distributed`` initialised every rank owns a contiguous run of the time slots of
        every file.  Sets ``_chunks`` = [(g_start, g_stop, my0, my1, local0)] and
        returns the rank's (start, stop) global time ranges; the rank's LOCAL time index runs
        over the concatenation of those ranges."""
        rank, world, _ = _dist_info()
        chunks, ranges, g_start, local = [], [], 0, 0
        for g_stop in time_chunks(times, self.max_times_per_file):
            my0, my1 = split_range(g_start, g_stop, world, rank)
            chunks.append((g_start, g_stop, my0, my1, local))
            ranges.append((my0, my1))
            local += my1 - my0
            g_start = g_stop
        self._chunks = chunks
        return ranges

    def _load_rows(self, soltab_name):
        """(Solutions restricted to this rank's time rows, full time axis)."""
        if self.solutions is not None:
            sol = self.solutions[soltab_name] if isinstance(self.solutions, dict) else self.solutions
            times = np.asarray(sol.time, np.float64)
            ranges = self._plan_rows(times)
            return (sol if _dist_info()[1] == 1 else sol.rows(ranges)), times
        from .h5parm_io import read_solutions, read_time_axis
        times = read_time_axis(self.input_h5parm_filename, self.input_solset_name, soltab_name)
        ranges = self._plan_rows(times)
        if _dist_info()[1] == 1:
            ranges = None
        return read_solutions(self.input_h5parm_filename, self.input_solset_name, soltab_name,
                              time_ranges=ranges), times

    def _local_rows(self):
        """Global time index of every local row."""
        return np.concatenate([np.arange(c[2], c[3]) for c in self._chunks]) if self._chunks else np.arange(0)

    @staticmethod
    def _all_reduce(array, op="sum"):
        """Element-wise reduction of a small host array over the ranks (metadata only)."""
        _, world, dist = _dist_info()
        if world == 1:
            return array
        t = torch.as_tensor(np.ascontiguousarray(array))
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN}[op])
        return t.cpu().numpy()

    # ------------------------------------------------------------------ write
    def slab_columns(self, cellsize_deg, planes=4):
        """(time, freq, antenna) columns per device slab.  A slab is a run of whole columns, not
        of whole time slots: one SKA-Low time slot (32 freqs x 512 stations x 2048^2) is 1.1 TB."""
        n_y, n_x = self.image_shape(cellsize_deg)
        return max(1, int(self.slab_bytes // (4 * planes * n_y * n_x)))

    def _out_format(self, smooth_pix):
        """(out_flags the slab kernels of this run fuse into their stores, planes stored per
        column).  Subclasses widen this; the base asks for native 4-plane slabs and a separate
        byte-swap pass."""
        return 0, 4

    def _blocks(self, local0, local1):
        """Yields ``(r0, r1, cols, amp)`` for runs of this rank's local time rows [r0, r1) in
        [local0, local1): cols = device float64 [(r1 - r0) * n_freq * n_ant, n_dir] of the values
        the cube kernels consume (whitened screens / referenced phases), amp = the matching
        [.., n_dir, 2] amplitudes of gain solutions or None.  The base serves them from the
        device-resident result of an eager fit(); KLScreen overrides it with the streamed fit."""
        per_t = len(self.freqs_ph) * len(self.station_names)
        n_d = self._dev_vals.shape[-1]
        rows = max(1, int(self.block_columns // per_t))
        for r0 in range(local0, local1, rows):
            r1 = min(local1, r0 + rows)
            cols = self._dev_vals[r0:r1].reshape(-1, n_d)
            amp = None if self.phase_only else self._dev_amp[r0:r1].reshape(-1, n_d, 2)
            yield r0, r1, cols, amp

    def write(self, out_dir, cellsize_deg, smooth_pix=0, ncpu=0):
        """Writes ``{name}_{g}.fits`` per gap-free time chunk and ``{name}.txt``
        (screen.py:260-394).

        Pipeline per device slab (a run of (time, freq, antenna) columns): the cube kernels write
        it in FITS byte order (and, for phase-only solutions, as plane pairs) on the main stream;
        a copy stream moves it to one of ``host_buffers`` pinned staging buffers as soon as it is
        complete, while the next slab is computed into the other device buffer; a flusher thread
        places each staged slab at its byte offset of the output file with ``write_threads``
        parallel gathered writes.  KL screens fit the next block of time slots on a third stream
        meanwhile (KLScreen._blocks)."""
        self.ncpu = ncpu
        t_write = time.perf_counter()
        rank, world, dist = _dist_info()
        eng = self.engine
        cuda = eng.device.type == "cuda"
        n_y, n_x = self.image_shape(cellsize_deg)
        npix = n_y * n_x
        n_f, n_a = len(self.freqs_ph), len(self.station_names)
        per_t = n_f * n_a
        times = np.asarray(self.times_ph, np.float64)
        if self._chunks is None:
            self._plan_rows(times)
        t0 = time.perf_counter()
        self._prepare(cellsize_deg, out_dir)
        if cuda:
            torch.cuda.synchronize(eng.device)
        self.timings["prepare_s"] = time.perf_counter() - t0
        out_flags, planes = self._out_format(smooth_pix)
        fused_swap = bool(out_flags & _lib.OUT_BIG_ENDIAN)
        # planes per column in the FILE: 4 for a-term ("gain") cubes, 1 for dTEC cubes, which have
        # no MATRIX axis (processing_utils.py:187-190)
        file_planes = 1 if self.aterm_type == "tec" else 4
        step = self.slab_columns(cellsize_deg, planes)
        # never allocate more than the rank's largest chunk needs
        most = max([(c[3] - c[2]) * per_t for c in self._chunks] + [1])
        step = min(step, most)
        slab_elems = step * planes * npix
        n_host = max(2, int(self.host_buffers))
        if self._rings is None or self._rings[0] != (slab_elems, n_host):
            # two device slabs and the pinned staging buffers, kept across write() calls (pinning
            # a GiB of host memory costs more than filling it)
            self._rings = None
            self._rings = ((slab_elems, n_host), [eng.empty((slab_elems,), torch.float32) for _ in range(2)],
                           [torch.empty(slab_elems, dtype=torch.float32, pin_memory=cuda) for _ in range(n_host)])
        _, dev_ring, host_ring = self._rings
        host_free = queue.Queue()
        for i in range(n_host):
            host_free.put(i)
        main = torch.cuda.current_stream(eng.device) if cuda else None
        if cuda and self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(eng.device)
        copied = [None, None]
        stats = {"pwrite_s": 0.0, "file_bytes": 0, "d2h_bytes": 0, "slabs": 0, "flush_wait_s": 0.0}
        flush_q = queue.Queue()
        failure = []
        pool = ThreadPoolExecutor(max(1, int(self.write_threads))) if self.sink == "file" else None
        n_place = max(1, int(self.write_threads))

        def place(dst, src, n_columns, src_bytes, copies):
            # multi-threaded copies into the shared mapping of the file (ssf_host_place; the call
            # releases the GIL, so the main thread keeps issuing slabs meanwhile)
            eng._call("ssf_host_place", dst, src, n_columns, src_bytes, copies, n_place)

        def flusher():
            while True:
                item = flush_q.get()
                try:
                    if item is None:
                        return
                    writer, col_rel, hb, ncol, event = item
                    try:
                        if event is not None:
                            event.synchronize()
                        if writer is not None and not failure:
                            data = host_ring[hb][:ncol * planes * npix].numpy().view(">f4")
                            stats["pwrite_s"] += writer.write_columns(col_rel, data, already_big_endian=True,
                                                                      pairs=planes == 2, pool=pool,
                                                                      place=place if self.file_mapping else None)
                            stats["file_bytes"] += ncol * 4 * file_planes * npix
                    except BaseException as exc:  # noqa: BLE001 - surfaced by the caller
                        failure.append(exc)
                    finally:
                        host_free.put(hb)
                finally:
                    flush_q.task_done()

        thread = threading.Thread(target=flusher, name="ssf-flush", daemon=True)
        thread.start()
        outfiles = []
        voxels = 0
        k = 0
        try:
            for gnum, (g_start, g_stop, my0, my1, local0) in enumerate(self._chunks):
                outfile = os.path.join(out_dir, f"{self.name}_{gnum}.fits")
                outfiles.append(outfile)
                writer = None
                if self.sink == "file":
                    shape = (g_stop - g_start, n_f, n_a) + ((4,) if file_planes == 4 else ()) + (n_y, n_x)
                    cards = fits.aterm_cards(shape, self.rad, self.dec, cellsize_deg, self.freqs_ph,
                                             times[g_start:g_stop], aterm_type=self.aterm_type)
                    if rank == 0:
                        writer = fits.CubeWriter(outfile, shape, cards)
                    if dist is not None:
                        dist.barrier()
                    if rank != 0:
                        writer = fits.CubeWriter(outfile, shape, cards, create=False)
                    if self.file_mapping and self.preallocate and my1 > my0:
                        writer.preallocate((my0 - g_start) * per_t, (my1 - my0) * per_t)
                for r0, _r1, cols_all, amp_all in self._blocks(local0, local0 + (my1 - my0)):
                    col_base = (my0 - g_start + (r0 - local0)) * per_t     # first column of the block in the file
                    for c0 in range(0, cols_all.shape[0], step):
                        if failure:
                            raise failure[0]
                        c1 = min(cols_all.shape[0], c0 + step)
                        ncol = c1 - c0
                        b = k & 1
                        if copied[b] is not None:
                            main.wait_event(copied[b])           # slab b has left the device
                        slab = dev_ring[b][:ncol * planes * npix].view(ncol, planes, n_y, n_x)
                        self._fill_columns(cols_all[c0:c1], slab, cellsize_deg, out_dir, smooth_pix, fix_nan=True,
                                           amp=None if amp_all is None else amp_all[c0:c1], out_flags=out_flags)
                        if not fused_swap:
                            eng.bswap32(slab)
                        t_wait = time.perf_counter()
                        hb = host_free.get()                     # a staging buffer the flusher is done with
                        stats["flush_wait_s"] += time.perf_counter() - t_wait
                        nbytes = ncol * planes * npix * 4
                        event = None
                        if cuda:
                            done = torch.cuda.Event()
                            done.record(main)
                            with torch.cuda.stream(self._copy_stream):
                                self._copy_stream.wait_event(done)
                                eng.copy_rows_d2h(host_ring[hb], nbytes, slab, nbytes, nbytes, 1)
                                event = torch.cuda.Event()
                                event.record(self._copy_stream)
                            copied[b] = event
                        else:
                            host_ring[hb][:ncol * planes * npix].copy_(slab.view(-1))
                        flush_q.put((writer, col_base + c0, hb, ncol, event))
                        stats["d2h_bytes"] += nbytes
                        stats["slabs"] += 1
                        voxels += ncol * npix
                        k += 1
                flush_q.join()
                if writer is not None:
                    writer.close()
                if failure:
                    raise failure[0]
        finally:
            flush_q.put(None)
            thread.join(timeout=60)
            if pool is not None:
                pool.shutdown(wait=True)
        if cuda:
            torch.cuda.synchronize(eng.device)
        eng.check_fit_status()                                   # deferred checks of the streamed fit
        if getattr(self, "write_soltabs", False):
            self.write_back()
        if dist is not None:
            dist.barrier()
        if rank == 0 and self.sink == "file":
            with open(os.path.join(out_dir, f"{self.name}.txt"), "w", encoding="utf8") as handle:
                handle.writelines([o + "\n" for o in outfiles])
        stats.update(voxels_written=voxels, write_s=time.perf_counter() - t_write, planes_copied=planes,
                     out_flags=out_flags, columns_per_slab=step, host_buffers=n_host,
                     write_threads=int(self.write_threads) if pool is not None else 0, sink=self.sink)
        self.timings.update(stats)
        return outfiles

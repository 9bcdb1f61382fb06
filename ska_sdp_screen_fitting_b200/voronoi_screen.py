"""
Voronoi-tessellated screens -- drop-in for the reference's ``voronoi_screen.py:24-351``.

``fit()`` re-references the phases on the device; the cube is produced per time slab by
``ssf_voronoi_lut`` (FP64 cos/sin of each solution, rounded once to float32) and
``ssf_voronoi_fill`` (template lookup, four coalesced plane stores per pixel), followed --
when smoothing is requested -- by the separable Gaussian of ``ssf_gauss_smooth`` and the
NaN fix, or by the algebraically equivalent contraction against pre-smoothed cell indicator
images (``smooth_mode = "linear"``).
"""
import os

import numpy as np
import torch

from . import _lib, fits, skymodel, template, wcs
from .kl_screen import reference_station
from .screen import Screen, _dist_info


class VoronoiScreen(Screen):
    """Class for Voronoi screens (reference: voronoi_screen.py:24-55)."""

    def __init__(self, name, h5parm_filename, skymodel_filename, rad, dec, width_ra, width_dec,
                 solset_name="sol000", phase_soltab_name="phase000", amplitude_soltab_name=None):
        super().__init__(name, h5parm_filename, skymodel_filename, rad, dec, width_ra, width_dec,
                         solset_name=solset_name, phase_soltab_name=phase_soltab_name,
                         amplitude_soltab_name=amplitude_soltab_name)
        self.data_rasertize_template = None
        self.polygons = None
        self.template_mode = "reference"      # "reference" (host pipeline) | "nearest" (CUDA kernel)
        # "direct" (fill -> gaussian_filter, the reference's order of operations), "linear"
        # (smoothed cell indicators contracted against the cos/sin LUT: the same result by
        # linearity at 16 B/voxel for any sigma), or "auto": linear unless the phases hold NaN
        # (0 * NaN would carry a NaN beyond the filter radius in the linear form)
        self.smooth_mode = "auto"
        self._has_nan = None                  # set by fit(); None = look at every slab
        self._amp_finite = None               # gain solutions: all amplitudes finite (looked at once)
        self.patch_radec = None               # optional [ndir, 2] deg override of the sky model
        self._tmpl_dev = None
        self._tmpl_key = None
        self._ind_dev = None
        self._ind_key = None

    def fit(self):
        """voronoi_screen.py:57-102: no fitting, the phases are referenced to one station."""
        sol, times = self._load_rows(self.input_phase_soltab_name)
        if sol.val.ndim != 4:
            raise NotImplementedError("phase soltabs with a pol axis are not supported (scalarphase only, "
                                      "as in the reference: voronoi_screen.py:69-70)")
        eng = self.engine
        # the reference station is global over time (processing_utils.py:560-572): per-antenna
        # weight sums of this rank's rows, summed over the ranks
        sums = self._all_reduce(np.sum(np.asarray(sol.weight), axis=(0, 1, 3), dtype=float), "sum")[:10]
        ref_ind = int(np.nonzero(sums == sums.max())[0][0])
        val = eng.tensor(sol.val, torch.float64)
        self._dev_vals = (val - val[:, :, ref_ind:ref_ind + 1, :]).contiguous()
        has_nan = np.array([int(np.isnan(np.asarray(sol.val)).any())], np.int32)
        self._has_nan = bool(-self._all_reduce(-has_nan, "min")[0])          # max over the ranks
        self.vals_ph = self._dev_vals.cpu().numpy()
        self.times_ph = times
        self.freqs_ph = sol.freq
        if self.phase_only:
            self.vals_amp = np.ones_like(self.vals_ph)
            self.times_amp = self.times_ph
            self.freqs_amp = self.freqs_ph
        else:
            amp = self._load_solutions(self.input_amplitude_soltab_name)      # voronoi_screen.py:84-88
            if amp.val.ndim != 5 or amp.val.shape[-1] < 2:
                raise ValueError("amplitude soltab must be [time, freq, ant, dir, pol] with XX and YY")
            self.log_amps = False
            self.vals_amp = np.array(amp.val, dtype=np.float64)
            self.times_amp = amp.time
            self.freqs_amp = amp.freq
        self.source_names = sol.dir
        self.source_dict = sol.source_dict
        self.source_positions = sol.source_positions
        self.station_names = sol.ant
        self.station_dict = sol.station_dict
        self.station_positions = sol.station_positions

    # ------------------------------------------------------------------ template
    def _patch_radec(self):
        if self.patch_radec is not None:
            arr = np.asarray(self.patch_radec, np.float64)
            return arr[:, 0], arr[:, 1]
        positions = skymodel.patch_positions(self.input_skymodel_filename)   # voronoi_screen.py:241-249
        radec = np.array([positions[name.strip("[]")] for name in self.source_names])
        return radec[:, 0], radec[:, 1]

    def make_rasertize_template(self, cellsize_deg, out_dir):
        """voronoi_screen.py:218-351.  Also writes ``{name}_template.fits`` (a one-time-slot
        zero cube, :229-232) when ``out_dir`` is given, because the reference leaves it behind
        and its test asserts it exists."""
        eng = self.engine
        n_y, n_x = self.image_shape(cellsize_deg)
        ra_deg, dec_deg = self._patch_radec()
        if out_dir is not None and _dist_info()[0] == 0:     # one writer: rank 0
            shape = (1, len(self.freqs_ph), len(self.station_names), 4, n_y, n_x)
            cards = fits.aterm_cards(shape, self.rad, self.dec, cellsize_deg, self.freqs_ph,
                                     np.asarray(self.times_ph)[0:1])
            fits.CubeWriter(os.path.join(out_dir, f"{self.name}_template.fits"), shape, cards).close()
        if self.template_mode == "reference":
            tmpl, _, _ = template.reference_template(ra_deg, dec_deg, self.rad, self.dec,
                                                     self.width_ra, cellsize_deg)
            self._tmpl_dev = eng.tensor(tmpl, torch.int32)
        elif self.template_mode == "nearest":
            px, py = wcs.patch_pixels(ra_deg, dec_deg, self.rad, self.dec, n_x, cellsize_deg)
            self._tmpl_dev = eng.voronoi_assign(eng.tensor(px, torch.float64),
                                                eng.tensor(py, torch.float64), n_x, n_y)
            tmpl = self._tmpl_dev.cpu().numpy()
        else:
            raise ValueError(f"unknown template_mode {self.template_mode!r}")
        self.data_rasertize_template = tmpl.astype(np.float64)
        self._ind_dev = None

    def _prepare(self, cellsize_deg, out_dir):
        key = (float(cellsize_deg), float(self.width_ra), float(self.rad), float(self.dec),
               self.template_mode)
        if self._tmpl_dev is None or self._tmpl_key != key:
            self.make_rasertize_template(cellsize_deg, out_dir)
            self._tmpl_key = key

    def _smoothed_indicators(self, smooth_pix):
        key = (self._tmpl_key, float(smooth_pix))
        if self._ind_dev is None or self._ind_key != key:
            eng = self.engine
            ind = eng.voronoi_indicators(self._tmpl_dev, len(self.source_names))
            self._ind_dev = eng.gauss_smooth(ind, smooth_pix)
            self._ind_key = key
        return self._ind_dev

    # ------------------------------------------------------------------ fill
    def _use_linear(self, smooth_pix, cols=None):
        """Whether a smoothed slab takes the linear (contraction) form."""
        if not (bool(smooth_pix) and smooth_pix > 0):
            return False
        if self.smooth_mode != "auto":
            return self.smooth_mode == "linear"
        has_nan = self._has_nan
        if has_nan is None:
            has_nan = cols is None or bool(torch.isnan(cols).any().item())
        if not self.phase_only and not has_nan:
            # gain solutions: a non-finite amplitude (NaN, or inf from a zero in log space)
            # would spread through the linear form just like a NaN phase; looked at once per run
            if self._amp_finite is None:
                self._amp_finite = self._dev_amp is not None and bool(torch.isfinite(self._dev_amp).all().item())
            has_nan = not self._amp_finite
        # tiny images (the repo's 17 x 17 test geometry) stay on the reference's own order of
        # operations: nothing to gain there, and it is the tighter parity
        return (not has_nan) and self._tmpl_dev.numel() >= 128 * 128

    def _out_format(self, smooth_pix):
        # the fill and the linear smoothed form store final values: FITS byte order, and plane
        # pairs for phase-only solutions (planes 2, 3 = planes 0, 1, voronoi_screen.py:194-198).
        # The direct smoothing path filters native 4-plane slabs in place and swaps afterwards.
        smooth = bool(smooth_pix) and smooth_pix > 0
        if smooth and not self._use_linear(smooth_pix):
            return 0, 4
        if self.phase_only:
            return _lib.OUT_BIG_ENDIAN | _lib.OUT_PLANES2, 2
        return _lib.OUT_BIG_ENDIAN, 4

    def interpolate(self, interp_kind="nearest"):
        super().interpolate(interp_kind)
        self._amp_finite = None

    def _fill_columns(self, cols, slab, cellsize_deg, out_dir, smooth_pix, fix_nan, amp=None, out_flags=0):
        self._prepare(cellsize_deg, out_dir)
        eng = self.engine
        npix = self._tmpl_dev.numel()
        smooth = bool(smooth_pix) and smooth_pix > 0
        if amp is not None:
            # gain solutions: XX / YY amplitudes on planes 0-1 / 2-3 (voronoi_screen.py:177-214);
            # smoothing in the reference's own order of operations (screen.py:353-378)
            lut4 = eng.voronoi_lut_gain(cols.contiguous(), amp.contiguous())
            if smooth and self._use_linear(smooth_pix, cols):
                # smoothing is linear in each of the four planes.  Seen as 2 x ncol "columns" of
                # one plane pair each -- (c, XX) -> planes 0, 1 and (c, YY) -> planes 2, 3 of
                # column c -- a gain slab IS a plane-pair slab of the phase-only contraction, so
                # the same tcgen05 kernel serves it at the same 16 B/voxel instead of the 64 B of
                # filtering the slab in place
                ncol, n_d = cols.shape
                lut_pairs = lut4.view(ncol, n_d, 2, 2).permute(0, 2, 1, 3).contiguous().view(2 * ncol, n_d, 2)
                if out_flags & _lib.OUT_PLANES2:
                    raise ValueError("gain slabs have four different planes")
                eng.voronoi_contract(self._smoothed_indicators(smooth_pix), lut_pairs, slab,
                                     out_flags=out_flags | _lib.OUT_PLANES2)
                return
            if smooth and out_flags:
                raise ValueError("the direct smoothing path needs native 4-plane slabs")
            eng.voronoi_fill_gain(self._tmpl_dev, lut4, slab, fix_nan=fix_nan and not smooth, out_flags=out_flags)
            if smooth:
                eng.gauss_smooth_slab(slab, smooth_pix, fix_nan=fix_nan)
            return
        lut = eng.voronoi_lut(cols.contiguous())
        if smooth and self._use_linear(smooth_pix, cols):
            eng.voronoi_contract(self._smoothed_indicators(smooth_pix), lut, slab, out_flags=out_flags)
        elif smooth:
            if out_flags:
                raise ValueError("the direct smoothing path needs native 4-plane slabs")
            eng.voronoi_fill(self._tmpl_dev, lut, slab, fix_nan=False)
            eng.gauss_smooth_slab(slab, smooth_pix, fix_nan=fix_nan)
        else:
            eng.voronoi_fill(self._tmpl_dev, lut, slab, fix_nan=fix_nan, out_flags=out_flags)

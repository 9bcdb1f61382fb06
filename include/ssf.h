/*
 * ssf.h -- C ABI of the B200-native screen-fitting hot path (libssf_b200.so).
 *
 * The reference (ska-sdp-screen-fitting @ 213c471) is pure Python; there is no FFI in it.
 * These entry points are what a binding for the hot path h5parm solutions -> a-term cube
 * replaces, function by function.  Each declaration cites the
 * reference code it stands in for.  Conventions:
 *
 *   - every function returns 0 on success, a negative SSF_E* code on failure;
 *     ssf_last_error() returns a thread-local message for the last failure
 *   - no exceptions, no torch types: plain pointers + sizes
 *   - pointers are DEVICE pointers unless the parameter name ends in _host
 *   - the caller owns every buffer; nothing is allocated behind the caller's back except
 *     by the *_host convenience entry points (which allocate, copy, run, copy back, free)
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *     all device entry points are asynchronous on it and re-entrant per (device, stream)
 *
 * Layouts (row-major, last index fastest):
 *   column  = one (time, freq, antenna) cell, flattened in that order (the h5parm `val`
 *             layout [time, freq, ant, dir] makes each column a contiguous ndir vector)
 *   slab    = float32 [ncol, 4, ny, nx]  -- the FITS cube layout [t, f, a, 4, y, x] of
 *             processing_utils.py:190-194 restricted to a run of columns;
 *             planes = (Re XX, Im XX, Re YY, Im YY), screen.py:319-324
 *   voxel   = one (pixel, column) cell = 4 float32 = 16 bytes
 */
#ifndef SSF_H
#define SSF_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSF_OK 0
#define SSF_EINVAL (-1)   /* bad argument (null pointer, size out of range) */
#define SSF_ECUDA (-2)    /* CUDA runtime error (message in ssf_last_error) */
#define SSF_ENOCONV (-3)  /* eigen-solver did not converge */
#define SSF_ENOTSUP (-4)  /* shape not supported by this kernel variant */

#define SSF_MAX_DIR 112   /* max directions (pierce points) per system (2 n^2 doubles of smem) */

/* Output format of the cube kernels (`out_flags`), fused into their store epilogues:
 *   SSF_OUT_BIG_ENDIAN  every float32 is stored byte-swapped, i.e. in the big-endian order of
 *                       a BITPIX -32 FITS cube (what astropy's hdu.writeto does on the host,
 *                       screen.py:381-382) -- no separate pass over the slab;
 *   SSF_OUT_PLANES2     only planes 0 (Re XX) and 1 (Im XX) are stored, slab = float32
 *                       [ncol, 2, ny, nx].  For phase-only solutions planes 2, 3 repeat planes
 *                       0, 1 (kl_screen.py:368-373, voronoi_screen.py:194-198); the FITS writer
 *                       places each pair twice, so half the bytes are written to HBM and cross
 *                       PCIe.  Not valid for the gain entry points (their planes differ). */
#define SSF_OUT_BIG_ENDIAN 1
#define SSF_OUT_PLANES2 2

int ssf_version(void);
const char *ssf_last_error(void);
/* Number of kernel launches issued through this library by the calling process so far. */
uint64_t ssf_launch_count(void);

/* ------------------------------------------------------------------------------------
 * KL setup.  Replaces _calculate_svd (stationscreen.py:390-430) for a BATCH of direction
 * subsets ("flag patterns"): for pattern p with unflagged set S_p (mask bit d of
 * masks[p*2 + d/64]) over the pierce points pp[ndir,2]:
 *     C   = -((|pi-pj|^2 / r0^2)^(beta/2)) / 2           on S_p
 *     U   = left singular vectors of C, singular values descending (|eigenvalues|)
 *     pinvC = pinv(C) with absolute cutoff `pinv_cut`    (scipy 1.7.3 rcond semantics)
 * Output matrices are [npat, ndir, ndir] float64, compacted to the top-left n_p x n_p
 * block (n_p = popcount(mask)), rows/cols in increasing direction order.  sv = [npat, ndir]
 * singular values.  One CTA per pattern, cyclic Jacobi in shared memory.
 */
int ssf_kl_decompose(const double *pp, int ndir, const uint64_t *masks, int npat,
                     double r0, double beta, double pinv_cut,
                     double *C, double *pinvC, double *U, double *sv, void *stream);

/* Device-resident cache of decompositions, so that a fit pass needs no host round trip
 * (the reference recomputes the SVD for every system with a flagged direction,
 * stationscreen.py:493-499; systems that share a flag pattern share the result here).
 * For the masks [nsys, 2] of one pass (ssf_kl_flag_outliers) it finds or creates the slot
 * of every system's pattern (pat [nsys], slot 0 = the full pattern, which the caller
 * decomposes into slot 0 beforehand with ssf_kl_decompose) and decomposes the slots that
 * are new, all stream-ordered:
 *   table_keys [hash_cap, 2] u64, table_slots [hash_cap] i32 (initialised to -1):
 *       open-addressing hash table, hash_cap = power of two >= 2 * max_slots
 *   slot_masks [max_slots, 2] u64; C, pinvC, U [max_slots, ndir, ndir], sv [max_slots, ndir]
 *   counters [4] i32 = {slots in use (start at 1), slots decomposed (start at 1),
 *       overflow flag, scratch}.  When a pass could run out of slots the cache forgets
 *       everything but slot 0 first; counters[2] = 1 only if one pass alone has more
 *       distinct patterns than max_slots - 1 (results are then invalid).
 *   status [1] i32: set to 1 if a Jacobi solve did not converge.
 */
int ssf_kl_patterns_update(const uint64_t *masks, int64_t nsys, int ndir, const double *pp,
                           double r0, double beta, double pinv_cut,
                           uint64_t *table_keys, int32_t *table_slots, int hash_cap,
                           uint64_t *slot_masks, int max_slots,
                           double *C, double *pinvC, double *U, double *sv,
                           int32_t *counters, int32_t *status, int32_t *pat, void *stream);

/* Grid width of ssf_kl_fit_pass in CTAs per SM, process-wide (0 = the default, 16).  The streamed
 * path (fits on a side stream beside the cube kernels) sets 2 so that fit CTAs fit next to the
 * resident evaluation CTAs instead of replacing them; an eager fit of a whole table wants 16. */
int ssf_kl_fit_ctas_per_sm(int ctas);

/* ------------------------------------------------------------------------------------
 * KL fit, one pass over all systems.  Replaces _fit_screen (stationscreen.py:433-594) and
 * the per-time body of _process_station (:681-779) for phase screens.
 *
 *   phase   [nsys, ndir] f64   referenced phases (rad)
 *   weights [nsys, ndir] f32   >0 = used
 *   pat     [nsys] i32         index of each system's flag pattern in the decompose output;
 *                              pattern 0 must be the FULL pattern (no direction flagged): its
 *                              pinv and its structure matrix serve the prediction at flagged
 *                              directions (:565-574) of every other pattern
 *   order   [nsys] f64 in/out  current screen order (the reference stores it as float)
 *   station_order [nsys] f64   per-system lower bound (= order of the station, :656)
 *   mode 0: iteration 0       -> one fit at min(order, n_unfl-1)         (:681-722)
 *   mode 1: iteration > 0     -> order-adjust state machine, first order iteration skips
 *                                the fit (reference quirk, :694-705), <= 3 refits (:692-779)
 *   mode + 2: amplitude type  -> `phase` holds amplitudes; log10 fit (:535-548), residual
 *                                amp - 10^fit (:576-590), log-residual chi^2 (:732-741)
 *   mode + 4: tec type        -> `phase` holds dTEC values, fitted as they are (:549-562),
 *                                residual val - fit, chi^2 = sum w resid^2 (:742-746)
 *   white, resid [nsys, ndir] f64 in/out (mode 1 reads the residual of the previous pass)
 * Systems with no unflagged direction are left untouched (:683-684).
 *   status  [1] i32 in/out     error flag of the pass (the caller zero-initialises it and
 *                              checks it afterwards); weights in (0, pinv_cut] are NOT an
 *                              error: such systems take the general pinv path (round 2).
 */
int ssf_kl_fit_pass(const double *phase, const float *weights, const int32_t *pat,
                    const double *pp, int ndir, int64_t nsys,
                    const double *C, const double *pinvC, const double *U,
                    double r0, double beta, double pinv_cut, int mode,
                    double *order, const double *station_order,
                    double *white, double *resid, int32_t *status, void *stream);

/* Replaces _flag_outliers (stationscreen.py:303-350) + normalize_phase / nancircstd
 * (processing_utils.py:73-132): per system, sigma = circular std of the unflagged
 * residuals over directions; weights of |resid| > nsigma*sigma are set to 0 in place.
 * Also emits each system's new 128-bit flag mask (masks_out [nsys, 2]).  */
int ssf_kl_flag_outliers(const double *resid, float *weights, int ndir, int64_t nsys,
                         double nsigma, uint64_t *masks_out, void *stream);

/* Amplitude-type flagging (_flag_outliers :341-348 on the log residual of :664-668):
 * systems are time-major with `ngroups` = freq*station*pol systems per time slot
 * (sys = t*ngroups + g); one weighted rms per group over all directions and times
 * (group_std [ngroups] scratch), weights of |log resid| > nsigma*rms set to 0 in place.
 * nsigma = inf flags nothing and only emits the masks.
 * val == NULL: tec type, the residual itself is clipped (:338-346 with :661-663). */
int ssf_kl_flag_outliers_amp(const double *val, const double *resid, float *weights, int ndir,
                             int64_t nsys, int64_t ngroups, double nsigma, double *group_std,
                             uint64_t *masks_out, void *stream);

/* ------------------------------------------------------------------------------------
 * KL evaluation.  Replaces KLScreen.make_matrix + calculate_kl_screen
 * (kl_screen.py:192-380, 411-449) for all columns of a slab at once.
 *
 * ssf_kl_basis: per-pixel interpolation basis (once per run)
 *     Bt[d, pix] = sum_e K(pix, e) * pinvC[e, d],  K(pix,e) = -((|g_pix - p_e|^2/r0^2)^(beta/2))/2
 *   with the separable grid g_pix = (x_coord[i], y_coord[j]), pix = j*nx + i
 *   (kl_screen.py:255-259, 444-448).  FP64 math, stored float32 direction-major
 *   [ndir, ny*nx] (pixel fastest: coalesced for SIMT, MN-major operand for tcgen05).
 * ssf_kl_coeffs: per-column coefficients  G[col, d] = sum_e C[d, e] * white[col, e]
 *   (FP64 math, stored float32 [ncol, ldg], zero padded); then
 *       Phi[pix, col] = Bt[:,pix] . G[col,:]  ==  K(pix,:) . white[col,:]   (kl_screen.py:449)
 *   because white = pinvC * fit lies in the range of pinvC*C.
 * ssf_kl_eval: slab[col, 0|2, pix] = cos Phi, slab[col, 1|3, pix] = sin Phi
 *   (kl_screen.py:368-373), NaN -> (1, 0) (screen.py:364-378) fused into the store.
 */
int ssf_kl_basis(const double *x_coord, const double *y_coord, int nx, int ny,
                 const double *pp, int ndir, const double *pinvC,
                 double r0, double beta, float *Bt, void *stream);
int ssf_kl_coeffs(const double *white, int64_t ncol, int ndir, const double *C,
                  float *G, int ldg, void *stream);
/* variant: 0 = FP32 SIMT, 1 = tcgen05 (3xTF32 split, FP32 accumulate in TMEM), -1 = auto */
int ssf_kl_eval(const float *Bt, const float *G, int ndir, int ldg, int64_t npix,
                int64_t ncol, float *slab, int variant, int out_flags, void *stream);

/* dTEC screens (aterm_type="tec", processing_utils.py:187-190: a 5-axis cube [t, f, a, y, x]
 * with no MATRIX axis): the same contraction without the cos/sin epilogue,
 *     slab[col, pix] = Phi[pix, col]        (float32 [ncol, ny*nx]; NaN kept)
 * out_flags: SSF_OUT_BIG_ENDIAN only. */
int ssf_kl_eval_tec(const float *Bt, const float *G, int ndir, int ldg, int64_t npix,
                    int64_t ncol, float *slab, int variant, int out_flags, void *stream);

/* Gain solutions (scalar phase + log10 XX/YY amplitude screens, kl_screen.py:319-378):
 * G3 [ncol, 3, ldg] f32 = coefficients (phase, Axx, Ayy) from ssf_kl_coeffs on the stacked
 * white vectors; slab[col] = (10^Axx cos Phi, 10^Axx sin Phi, 10^Ayy cos Phi, 10^Ayy sin Phi),
 * NaN -> (1, 0, 1, 0) fused into the store.  variant: -1 auto, 0 FP32 SIMT, 1 tcgen05
 * (16 columns x 3 accumulator groups per 64-wide tile; same kernel as ssf_kl_eval). */
int ssf_kl_eval_gain(const float *Bt, const float *G3, int ndir, int ldg, int64_t npix,
                     int64_t ncol, float *slab, int variant, int out_flags, void *stream);

/* ------------------------------------------------------------------------------------
 * Voronoi.  ssf_voronoi_assign: float64 nearest patch per integer pixel, lowest index
 * wins ties; tmpl[j*nx+i] = patch index + 1 (the values of data_rasertize_template,
 * voronoi_screen.py:329-350).  ssf_voronoi_lut: cos/sin of the referenced phases in FP64,
 * rounded once to float32: lut[col, d, 2] (voronoi_screen.py:203-214 + the float32 store
 * of screen.py:343).  ssf_voronoi_fill: slab[col, p, pix] = lut[col, tmpl[pix]-1, p&1]
 * (VoronoiScreen.make_matrix, voronoi_screen.py:169-216), NaN fix fused.
 */
int ssf_voronoi_assign(const double *px, const double *py, int ndir, int nx, int ny,
                       int32_t *tmpl, void *stream);
int ssf_voronoi_lut(const double *phase, const double *amp, int64_t ncol, int ndir,
                    float *lut, void *stream);
int ssf_voronoi_fill(const int32_t *tmpl, const float *lut, int ndir, int64_t npix,
                     int64_t ncol, float *slab, int fix_nan, int out_flags, void *stream);

/* Gain solutions (scalar phase + XX/YY amplitudes, voronoi_screen.py:177-214):
 * amp [ncol, ndir, 2] f64 = (XX, YY) amplitudes; lut4 [ncol, ndir, 4] f32 =
 * (axx cos, axx sin, ayy cos, ayy sin); slab[col, p, pix] = lut4[col, tmpl[pix]-1, p]. */
int ssf_voronoi_lut_gain(const double *phase, const double *amp, int64_t ncol, int ndir,
                         float *lut4, void *stream);
int ssf_voronoi_fill_gain(const int32_t *tmpl, const float *lut4, int ndir, int64_t npix,
                          int64_t ncol, float *slab, int fix_nan, int out_flags, void *stream);

/* Smoothed Voronoi through linearity: slab[col, p, pix] = sum_d S[d, pix] * lut[col, d, p&1]
 * where S[d, :] is the Gaussian-smoothed indicator image of cell d (built with
 * ssf_voronoi_indicators + ssf_gauss_smooth).  Same result as fill -> gaussian_filter
 * (screen.py:353-362) up to float32 rounding of the intermediate. */
int ssf_voronoi_indicators(const int32_t *tmpl, int ndir, int64_t npix, float *ind,
                           void *stream);
int ssf_voronoi_contract(const float *S, const float *lut, int ndir, int64_t npix,
                         int64_t ncol, float *slab, int variant, int out_flags, void *stream);

/* ------------------------------------------------------------------------------------
 * Separable Gaussian smoothing, in place, of nimg images [ny, nx] float32: y pass then x
 * pass, scipy 'reflect' boundary, radius int(4*sigma+0.5), weights exp(-x^2/(2 sigma^2))
 * normalised, float32 storage between the passes (ndimage.gaussian_filter with
 * sigma=(0,s,s), screen.py:358-362).  `scratch` holds nimg*ny*nx floats. */
int ssf_gauss_smooth(float *images, int64_t nimg, int ny, int nx, double sigma,
                     float *scratch, void *stream);

/* The same filter on the 4 * ncol planes of a slab [ncol, 4, ny, nx], with the NaN fix of
 * screen.py:364-378 (below) fused into the stores of the second pass when `fix_nan` is set: the
 * reference's order of operations for slabs that carry NaN (screen.py:353-378) in two passes
 * over the slab instead of three. */
int ssf_gauss_smooth_slab(float *slab, int64_t ncol, int ny, int nx, double sigma,
                          float *scratch, int fix_nan, void *stream);

/* ssf_gauss_smooth without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same filter, same kernels and the
 * same results as ssf_gauss_smooth, but the intermediate image and the taps live in a file-scope
 * device array inside the library (SSF_SMOOTH_STATIC_MAX_FLOATS floats = 16 MiB, plus the taps;
 * every GPU that loads the library holds it until the process exits), so the call makes no
 * cudaMalloc / cudaFree / stream-ordered allocation of any kind.
 *   limit   nimg * ny * nx <= SSF_SMOOTH_STATIC_MAX_FLOATS, else SSF_EINVAL with a message that
 *           names the size (ssf_gauss_smooth_static_max_floats() returns the built-in limit);
 *   sharing every call on a device uses the same array, which keeps the last call's contents
 *           (nothing in it needs to start at zero: each call rewrites all it reads).  Calls that
 *           overlap on the GPU -- two streams, or two host threads -- race on it: serialise them. */
#define SSF_SMOOTH_STATIC_MAX_FLOATS (4 * 1024 * 1024)
int ssf_gauss_smooth_static(float *images, int64_t nimg, int ny, int nx, double sigma,
                            void *stream);
int64_t ssf_gauss_smooth_static_max_floats(void);

/* NaN -> 1.0 (planes 0, 2) / 0.0 (planes 1, 3), screen.py:364-378, on a slab. */
int ssf_nan_fix(float *slab, int64_t ncol, int64_t npix, void *stream);

/* In-place byte swap of `count` 32-bit words: float32 little-endian -> the big-endian
 * storage order of a BITPIX -32 FITS cube (astropy does this inside hdu.writeto,
 * screen.py:381-382).  `data` must be 16-byte aligned. */
int ssf_bswap32(void *data, int64_t count, void *stream);

/* Strided device -> host copy of `rows` runs of `width_bytes` (cudaMemcpy2DAsync on `stream`):
 * how a slab of column pairs [ncol, 2, ny, nx] (SSF_OUT_PLANES2) lands at planes 0-1 of each
 * 4-plane column of the pinned host image of the FITS cube (src_pitch = 2 planes, dst_pitch =
 * 4 planes), or a slab of whole columns lands contiguously (both pitches = width).  The FITS
 * write of the reference is hdu.writeto (screen.py:381-382); this is the device half of it. */
int ssf_copy_rows_d2h(void *dst_host, int64_t dst_pitch, const void *src_dev, int64_t src_pitch,
                      int64_t width_bytes, int64_t rows, void *stream);

/* Host only: places a staged slab (pinned host memory, FITS byte order) into the memory-mapped
 * FITS cube.  Column c of `src` (`src_bytes` bytes) is copied to dst + c * copies * src_bytes,
 * `copies` times back to back: copies = 1 for whole columns, 2 for the plane pairs of a
 * phase-only cube (SSF_OUT_PLANES2), whose planes 2, 3 repeat planes 0, 1.  `threads` worker
 * threads share the bytes.  This is the host half of the reference's hdu.writeto
 * (screen.py:381-382); pwrite calls on one file serialise on the inode lock, copies into a
 * shared mapping do not. */
int ssf_host_place(void *dst, const void *src, int64_t ncol, int64_t src_bytes, int copies, int threads);

/* ------------------------------------------------------------------------------------
 * Host convenience entry points (what a ctypes/cffi binding of the reference would call):
 * HOST buffers in, HOST slab out; allocation, H2D, kernels, D2H inside.
 */
int ssf_voronoi_cube_host(const int32_t *tmpl_host, const double *phase_host, int ndir,
                          int ny, int nx, int64_t ncol, double smooth_pix,
                          float *slab_host);
int ssf_kl_cube_host(const double *x_host, const double *y_host, int nx, int ny,
                     const double *pp_host, int ndir, const double *white_host,
                     int64_t ncol, double r0, double beta, double pinv_cut,
                     float *slab_host);

#ifdef __cplusplus
}
#endif
#endif /* SSF_H */

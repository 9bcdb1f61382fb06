"""
TEST INFRASTRUCTURE (oracle.port): LOOP-FAITHFUL restatement of the reference's CPU hot
loops, used only to time the CPU baseline (``bench.py`` ``cpu_baseline`` / ``--impl
reference``).  The reference is pure Python and cannot travel to the GPU box, so this module
restates its loop STRUCTURE -- not just its arithmetic -- one numpy call per reference numpy
call, with the same multiprocessing layout:

* ``kl_make_matrix``: a ``multiprocessing.Pool`` of ``ncpu`` workers mapping over time slots;
  each task runs the per-pixel double loop of ``calculate_kl_screen`` (kl_screen.py:294-317,
  :411-449); inputs reach the workers as fork-inherited module globals, output through a
  shared ``RawArray`` -- as in the reference.
* ``voronoi_make_matrix`` + ``smooth_and_fix``: the serial per-polygon / per-time fill of
  voronoi_screen.py:177-214 and the per-time ``gaussian_filter`` + NaN fix of
  screen.py:353-378.

``oracle.port.screens`` holds the vectorised arithmetic used for parity; this file is about
the COST of the reference's implementation.  Results are checked equal in tests/test_oracle.
"""
import itertools
import multiprocessing
from multiprocessing import Pool, RawArray

import numpy as np
from scipy import ndimage

SCREEN_PH = None
PIERCEPOINTS = None
X_COORD = None
Y_COORD = None
VAR_DICT = {}


def _init_worker(shared_val, val_shape):
    VAR_DICT["shared_val"] = shared_val
    VAR_DICT["val_shape"] = val_shape


def _calculate_kl_screen(args):
    """kl_screen.py:411-449, statement for statement."""
    k, n_piercepoints, beta_val, r_0 = args
    tmp = np.frombuffer(VAR_DICT["shared_val"], dtype=np.float64).reshape(VAR_DICT["val_shape"])
    f_matrix = SCREEN_PH[:, k].reshape(n_piercepoints)
    for i, x_i in enumerate(X_COORD):
        for j, y_i in enumerate(Y_COORD):
            p_array = np.array([x_i, y_i, 0.0])
            d_square = np.sum(np.square(PIERCEPOINTS - p_array), axis=1)
            c_matrix = -((d_square / (r_0 ** 2)) ** (beta_val / 2.0)) / 2.0
            tmp[i, j, k] = np.dot(c_matrix, f_matrix)


def kl_make_matrix(x_coord, y_coord, piercepoints, white_td, ncpu=0, beta_val=5.0 / 3.0, r_0=100.0):
    """KLScreen.make_matrix (kl_screen.py:192-380, phase-only) for one (freq, station):
    white_td [n_times, n_dir] -> float64 [n_times, 4, ny, nx].  A new Pool per call, like the
    reference (screen.py:335-351 calls make_matrix per (freq, station))."""
    global SCREEN_PH, PIERCEPOINTS, X_COORD, Y_COORD
    SCREEN_PH = np.ascontiguousarray(np.asarray(white_td, np.float64).T)       # [dir, time]
    PIERCEPOINTS = np.asarray(piercepoints, np.float64)
    X_COORD = np.asarray(x_coord, np.float64)
    Y_COORD = np.asarray(y_coord, np.float64)
    n_times = SCREEN_PH.shape[1]
    if ncpu == 0:
        ncpu = multiprocessing.cpu_count()
    val_shape = (len(X_COORD), len(Y_COORD), n_times)
    shared_val = RawArray("d", int(val_shape[0] * val_shape[1] * val_shape[2]))
    ctx = multiprocessing.get_context("fork")
    with ctx.Pool(processes=ncpu, initializer=_init_worker, initargs=(shared_val, val_shape)) as pool:
        pool.map(_calculate_kl_screen,
                 zip(range(n_times), itertools.repeat(PIERCEPOINTS.shape[0]),
                     itertools.repeat(beta_val), itertools.repeat(r_0)))
    val_phase = np.frombuffer(shared_val, dtype=np.float64).reshape(val_shape).copy()
    data = np.zeros((n_times, 4, len(Y_COORD), len(X_COORD)))
    for t in range(n_times):                                                  # :366-378
        data[t, 0, :, :] = np.cos(val_phase[:, :, t].T)
        data[t, 2, :, :] = data[t, 0, :, :]
        data[t, 1, :, :] = np.sin(val_phase[:, :, t].T)
        data[t, 3, :, :] = data[t, 1, :, :]
    return data


def voronoi_make_matrix(template, phases_td):
    """VoronoiScreen.make_matrix (voronoi_screen.py:169-216), phase-only (amplitudes are ones)."""
    n_times, n_dir = phases_td.shape
    data = np.zeros((n_times, 4) + template.shape)
    for d in range(n_dir):
        ind = np.where(template == d + 1)
        if not len(ind[0]):
            continue
        val_phase = phases_td[:, d]
        val_amp = np.ones(n_times)
        for t in range(n_times):
            data[t, 0, ind[0], ind[1]] = val_amp[t] * np.cos(val_phase[t])
            data[t, 2, ind[0], ind[1]] = val_amp[t] * np.cos(val_phase[t])
            data[t, 1, ind[0], ind[1]] = val_amp[t] * np.sin(val_phase[t])
            data[t, 3, ind[0], ind[1]] = val_amp[t] * np.sin(val_phase[t])
    return data


def smooth_and_fix(cube_f32, smooth_pix):
    """screen.py:353-378 on a float32 [nt, 4, ny, nx] block of one (freq, station)."""
    if smooth_pix > 0:
        for t in range(cube_f32.shape[0]):
            cube_f32[t] = ndimage.gaussian_filter(cube_f32[t], sigma=(0, smooth_pix, smooth_pix), order=0)
    for t in range(cube_f32.shape[0]):
        for p_val in range(4):
            nanval = 0.0 if p_val % 2 else 1.0
            cube_f32[t, p_val][np.isnan(cube_f32[t, p_val])] = nanval
    return cube_f32

"""ssf_gauss_smooth_static -- the smoothing entry point that allocates nothing (its scratch is a
fixed device array inside the library, include/ssf.h) -- against ssf_gauss_smooth, which takes
caller-owned scratch.  Both launch the same kernels with the same taps, so the results are
expected to be identical; the assertion uses the smoothing tolerance of the other tests.

Collected last on purpose: a failure here hides nothing.
The calls go through the C-ABI with ctypes; torch only owns the input / output device memory,
allocated BEFORE the calls under test.
"""
import ctypes

import numpy as np
import pytest
import torch

import parity_checks as pc


def _smooth_both(eng, shape, sigma, seed):
    """(static result, ssf_gauss_smooth result) for one random stack of images."""
    lib = eng._lib
    rng = np.random.default_rng(seed)
    host = rng.normal(size=shape).astype(np.float32)
    a = eng.tensor(host, torch.float32)
    b = eng.tensor(host, torch.float32)
    scratch = torch.empty_like(b)
    ny, nx = shape[-2:]
    nimg = host.size // (ny * nx)
    stream = ctypes.c_void_p(eng.stream())
    args = (ctypes.c_int64(nimg), ctypes.c_int(ny), ctypes.c_int(nx), ctypes.c_double(sigma))
    rc = lib.ssf_gauss_smooth_static(ctypes.c_void_p(a.data_ptr()), *args, stream)
    assert rc == 0, lib.ssf_last_error()
    rc = lib.ssf_gauss_smooth(ctypes.c_void_p(b.data_ptr()), *args, ctypes.c_void_p(scratch.data_ptr()), stream)
    assert rc == 0
    return a.cpu().numpy(), b.cpu().numpy(), host


def _check_static_smoothing(eng):
    lib = eng._lib
    limit = int(lib.ssf_gauss_smooth_static_max_floats())
    assert limit == 4 * 1024 * 1024
    # two calls in a row on the shared scratch: the second is smaller, has another radius and must
    # not see anything of the first; then the tap-per-output kernels (radius > 335), then a stack
    # that fills the array to the last float
    cases = [((7, 96, 131), 3.0, 11), ((3, 33, 20), 0.8, 12), ((2, 40, 50), 90.0, 13), ((4, 1024, 1024), 2.0, 14)]
    for shape, sigma, seed in cases:
        got, want, host = _smooth_both(eng, shape, sigma, seed)
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= pc.TOL_SMOOTH * max(1.0, float(np.abs(host).max())), (shape, sigma)
    # the first case again: same answer as the first time (nothing left behind by the calls between)
    again, want, _ = _smooth_both(eng, *cases[0])
    assert np.array_equal(again, want)
    # and against the oracle, like check_smooth does for the allocating entry point
    from oracle.port import screens

    got, _, host = _smooth_both(eng, (3, 19, 37), 2.0, 2)
    assert np.abs(got - screens.smooth_block(host, 2.0)).max() < pc.TOL_SMOOTH * max(1.0, float(np.abs(host).max()))
    # one pixel beyond the limit: refused with a message that names the sizes, nothing launched
    too_many = eng.tensor(np.zeros((5, 1024, 1024), np.float32), torch.float32)
    before = too_many.clone()
    rc = lib.ssf_gauss_smooth_static(ctypes.c_void_p(too_many.data_ptr()), ctypes.c_int64(5), ctypes.c_int(1024),
                                     ctypes.c_int(1024), ctypes.c_double(2.0), ctypes.c_void_p(eng.stream()))
    assert rc != 0
    message = lib.ssf_last_error().decode()
    assert "static" in message and str(limit) in message and "5242880" in message
    assert torch.equal(too_many, before)
    with pytest.raises(Exception, match="static array holds"):
        eng.gauss_smooth_static(too_many, 2.0)


@pytest.mark.gpu
def test_static_scratch_smoothing_matches_the_allocating_entry_point(gpu_engine):
    _check_static_smoothing(gpu_engine)


@pytest.mark.gpu
def test_static_scratch_smoothing_allocates_nothing(gpu_engine):
    """Device memory in use (driver's view) does not move across calls of the new entry point."""
    eng = gpu_engine
    img = eng.tensor(np.random.default_rng(3).normal(size=(8, 256, 256)).astype(np.float32), torch.float32)
    eng.gauss_smooth_static(img, 2.0)                      # set-up call: symbol address looked up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    reserved0 = torch.cuda.memory_reserved()
    for sigma in (1.0, 2.5, 6.0):
        eng.gauss_smooth_static(img, sigma)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_reserved() == reserved0
    assert free1 >= free0


def test_static_scratch_smoothing_logic_on_the_emulator(emu_engine):
    _check_static_smoothing(emu_engine)

"""GPU debug helper: error of ssf_kl_eval variants 0/1 against the float64 oracle."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import parity_checks as pc
from oracle.port import klfit, screens, wcs as owcs
from ska_sdp_screen_fitting_b200.engine import Engine

eng = Engine(device="cuda:0")

def case(n, n_d, ncol, seed, cell=0.1):
    rng = np.random.default_rng(seed)
    pp, mid_ra, mid_dec = klfit.piercepoints(pc.make_directions(rng, n_d, spread=1.5))
    x, y = owcs.kl_grid_coords(126.23, 64.5, (n - 0.5) * cell, cell, mid_ra, mid_dec)
    c_o, p_o, _ = klfit.calculate_svd(pp)
    fit = rng.normal(0, 0.7, (ncol, n_d))
    white = fit @ p_o.T
    ppt = eng.tensor(pp[:, :2], torch.float64)
    full = torch.tensor([pc.mask_of(range(n_d))], dtype=torch.int64, device=eng.device)
    c, p, _, _ = eng.kl_decompose(ppt, full)
    bt = eng.kl_basis(eng.tensor(x, torch.float64), eng.tensor(y, torch.float64), ppt, p[0].contiguous())
    g = eng.kl_coeffs(eng.tensor(white, torch.float64), c[0].contiguous())
    want = screens.kl_make_matrix(x, y, pp, white)
    out = {}
    for variant in (0, 1):
        slab = eng.empty((ncol, 4, n, n), torch.float32)
        slab.fill_(7.0)
        eng.kl_eval(bt, g, n_d, slab, variant)
        got = slab.cpu().numpy()
        err = np.abs(got - want)
        out[variant] = got
        print(f"n={n} n_d={n_d} ncol={ncol} cell={cell} variant={variant}: max err {err.max():.3e} "
              f"mean {err.mean():.3e} |Bt|max {float(bt.abs().max()):.3g} "
              f"worst col {int(err.reshape(ncol, -1).max(1).argmax())} untouched {(got == 7.0).sum()}", flush=True)
    print(f"   variant 0 vs 1: {np.abs(out[0] - out[1]).max():.3e}")

case(64, 45, 70, 9)
case(128, 45, 700, 15)
case(128, 45, 700, 15, cell=0.03)
case(128, 45, 3840, 16, cell=0.03)
case(128, 100, 2600, 17, cell=0.03)

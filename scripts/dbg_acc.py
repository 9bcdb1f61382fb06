"""GPU debug helper: accumulation error of the contraction variants on random operands."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_screen_fitting_b200.engine import Engine
eng = Engine(device="cuda:0")

def tf32(a):
    b = a.view(np.uint32).astype(np.uint64) + 0x1000
    return (b & 0xFFFFE000).astype(np.uint32).view(np.float32)

def run(n_d, ncol, npix, pre_round, positive, seed=1):
    rng = np.random.default_rng(seed)
    s = rng.normal(0, 1, (n_d, npix)).astype(np.float32)
    l = rng.normal(0, 1, (ncol, n_d, 2)).astype(np.float32)
    if positive:
        s, l = np.abs(s), np.abs(l)
    if pre_round:
        s, l = tf32(s), tf32(l)
    want = l[:, :, 0].astype(np.float64) @ s.astype(np.float64)
    for variant in (0, 1):
        slab = eng.empty((ncol, 4, npix), torch.float32)
        eng.voronoi_contract(eng.tensor(s, torch.float32), eng.tensor(l, torch.float32), slab, variant)
        got = slab[:, 0].cpu().numpy().astype(np.float64)
        err = got - want
        ulp = np.abs(want) * 2.0 ** -24
        print(f"n_d={n_d} pre_round={pre_round} positive={positive} variant={variant}: "
              f"mean signed err/ulp {np.mean(err / ulp):+.3f}  rms err/ulp {np.sqrt(np.mean((err / ulp) ** 2)):.3f} "
              f"max |err| {np.abs(err).max():.3e}", flush=True)

for pos in (True, False):
    for pr in (True, False):
        run(48, 64, 4096, pr, pos)
run(104, 64, 4096, False, True)

"""GPU debug helper: contraction kernel time vs image size / variant / kind on random operands."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ska_sdp_screen_fitting_b200.engine import Engine
eng = Engine(device="cuda:0")

def run(kind, n, n_d, ncol, variant):
    npix = n * n
    bt = torch.randn(n_d, npix, device="cuda", dtype=torch.float32)
    slab = torch.empty(ncol, 4, npix, device="cuda", dtype=torch.float32)
    if kind == "kl":
        ldg = (n_d + 3) // 4 * 4
        g = torch.zeros(ncol, ldg, device="cuda", dtype=torch.float32)
        g[:, :n_d] = torch.randn(ncol, n_d, device="cuda")
        f = lambda: eng.kl_eval(bt, g, n_d, slab, variant)
    else:
        lut = torch.randn(ncol, n_d, 2, device="cuda", dtype=torch.float32)
        f = lambda: eng.voronoi_contract(bt, lut, slab, variant)
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); f(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ms = sorted(ts)[1]
    gbs = ncol * npix * 16 / ms / 1e6
    print(f"{kind:3s} n={n:5d} n_d={n_d:3d} ncol={ncol:5d} variant={variant}: {ms:7.3f} ms  {gbs:7.0f} GB/s  frac {gbs/6528.7:.3f}", flush=True)
    del slab, bt

for c in [("kl", 512, 45, 5952), ("kl", 512, 45, 1488), ("kl", 1024, 45, 1488), ("kl", 2048, 100, 372), ("kl", 2048, 45, 372),
          ("vo", 512, 45, 5952), ("vo", 1024, 45, 1488), ("kl", 512, 100, 5952), ("kl", 512, 9, 5952)]:
    for variant in (1, 0):
        run(*c, variant)

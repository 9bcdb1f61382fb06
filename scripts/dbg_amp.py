import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import parity_checks as pc
from ska_sdp_screen_fitting_b200.engine import Engine
eng = Engine(device="cuda:0")
orig = pc.klfit.run_amplitude
hold = {}
def spy(*a, **k):
    hold["want"] = orig(*a, **k); return hold["want"]
pc.klfit.run_amplitude = spy
import ska_sdp_screen_fitting_b200.kl_screen as ks
orig_fit = ks.KLScreen.fit_amplitude_device
def spy2(self, *a, **k):
    hold["got"] = orig_fit(self, *a, **k); return hold["got"]
ks.KLScreen.fit_amplitude_device = spy2
for kw in (dict(), dict(seed=29, n_t=9, n_f=3, n_a=6, n_d=45)):
    try:
        pc.check_fit_amplitude(eng, **kw); print("ok", kw)
    except AssertionError:
        w, r = hold["got"][0].cpu().numpy(), hold["got"][1].cpu().numpy()
        d = np.abs(r - hold["want"]["resid"]); print(kw, "nan got/want", np.isnan(r).sum(), np.isnan(hold["want"]["resid"]).sum())
        i = np.unravel_index(np.nanargmax(d), d.shape); print("max", np.nanmax(d), i, r[i], hold["want"]["resid"][i], "white diff", np.abs(w-hold["want"]["screen"]).max())
        print("n bad", (d > 1e-9).sum(), np.argwhere(d > 1e-9)[:10])

"""Where does the host time of one KLScreen.fit_device call go?  (GPU box; cProfile + wall clock)"""
import cProfile
import os
import pstats
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import synthetic_cases as synthetic
from ska_sdp_screen_fitting_b200.engine import Engine
from ska_sdp_screen_fitting_b200.kl_screen import KLScreen

slots = int(sys.argv[1]) if len(sys.argv) > 1 else 16
case = synthetic.make_case("lofar", seed=3, n_times=slots * 6)
eng = Engine(device="cuda:0")
kl = KLScreen("p", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
kl._engine = eng
kl.setup(case.solutions)
val = eng.tensor(case.solutions.val, torch.float64)
wt = eng.tensor(case.solutions.weight, torch.float32)
for rep in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    kl.fit_device(val[rep * slots:(rep + 1) * slots], wt[rep * slots:(rep + 1) * slots], check=False)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"rep {rep}: host {1e3 * (t1 - t0):.2f} ms, +drain {1e3 * (t2 - t1):.2f} ms, patterns {kl._patterns.stats()[0]}")
pr = cProfile.Profile()
pr.enable()
kl.fit_device(val[5 * slots:6 * slots], wt[5 * slots:6 * slots], check=False)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(25)

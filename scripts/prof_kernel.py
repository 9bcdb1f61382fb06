"""One cube kernel alone, on random operands, timed with CUDA events (and a float64 check of a
sample of voxels) -- the quick loop for kernel work, and the command ncu wraps.

    python scripts/prof_kernel.py MODE [--ndir 45] [--n 512] [--ncol 5952] [--reps 10] [--flags 0]
                                       [--variant -1] [--check]

MODE: kl | gain | voronoi | tec | fill | fillgain.  Prints one line:
    MODE ndir n ncol : ms per launch, GB/s (16 B/voxel x planes/4), fraction of the measured HBM peak
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("mode", choices=["kl", "gain", "voronoi", "tec", "fill", "fillgain"])
    ap.add_argument("--ndir", type=int, default=45)
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--ncol", type=int, default=5952)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--flags", type=int, default=0)
    ap.add_argument("--variant", type=int, default=-1)
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    from ska_sdp_screen_fitting_b200.engine import Engine

    eng = Engine(device="cuda:0")
    dev = eng.device
    gen = torch.Generator(device=dev)
    gen.manual_seed(5)
    nd, npix, ncol = args.ndir, args.n * args.n, args.ncol
    ldg = (nd + 3) // 4 * 4
    planes = 1 if args.mode == "tec" else (2 if args.flags & 2 else 4)
    ring = [torch.empty(ncol * planes * npix, dtype=torch.float32, device=dev) for _ in range(2)]
    # cardinal-basis-like operand: O(1) values that sum to one over the directions
    bt = torch.rand((nd, npix), generator=gen, device=dev, dtype=torch.float32)
    bt /= bt.sum(0, keepdim=True)
    bt = bt * 3.0 - 2.0 / nd
    if args.mode in ("kl", "tec"):
        g = torch.zeros((ncol, ldg), device=dev, dtype=torch.float32)
        g[:, :nd] = torch.randn((ncol, nd), generator=gen, device=dev) * 2.0

        def run(slab):
            fn = eng.kl_eval if args.mode == "kl" else eng.kl_eval_tec
            fn(bt, g, nd, slab, variant=args.variant, out_flags=args.flags)
    elif args.mode == "gain":
        g = torch.zeros((ncol, 3, ldg), device=dev, dtype=torch.float32)
        g[:, 0, :nd] = torch.randn((ncol, nd), generator=gen, device=dev) * 2.0
        g[:, 1:, :nd] = torch.randn((ncol, 2, nd), generator=gen, device=dev) * 0.1

        def run(slab):
            eng.kl_eval_gain(bt, g, nd, slab, variant=args.variant, out_flags=args.flags)
    elif args.mode == "voronoi":
        g = torch.randn((ncol, nd, 2), generator=gen, device=dev, dtype=torch.float32)
        ind = (bt.abs() / bt.abs().sum(0, keepdim=True)).contiguous()

        def run(slab):
            eng.voronoi_contract(ind, g, slab, variant=args.variant, out_flags=args.flags)
    else:
        tmpl = torch.randint(1, nd + 1, (args.n, args.n), generator=gen, device=dev, dtype=torch.int32)
        if args.mode == "fill":
            g = torch.randn((ncol, nd, 2), generator=gen, device=dev, dtype=torch.float32)

            def run(slab):
                eng.voronoi_fill(tmpl, g, slab, out_flags=args.flags)
        else:
            g = torch.randn((ncol, nd, 4), generator=gen, device=dev, dtype=torch.float32)

            def run(slab):
                eng.voronoi_fill_gain(tmpl, g, slab, out_flags=args.flags)
    for i in range(args.warmup):
        run(ring[i & 1])
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.reps)]
    for i, (a, b) in enumerate(ev):
        a.record()
        run(ring[i & 1])
        b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in ev)
    med = ms[len(ms) // 2]
    nbytes = ncol * npix * 4 * planes
    peak = 6555.2
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:  # noqa: BLE001
        pass
    gbs = nbytes / med / 1e6
    line = (f"{args.mode} ndir={nd} n={args.n} ncol={ncol} flags={args.flags} variant={args.variant}: "
            f"median {med:.3f} ms (min {ms[0]:.3f}), {gbs:.0f} GB/s, {gbs / peak:.3f} of measured HBM peak")
    if args.check and args.mode in ("kl", "gain", "tec") and not args.flags & 1:
        last = ring[(args.reps - 1) & 1].view(ncol, planes, npix)
        cols = torch.randint(0, ncol, (64,), generator=gen, device=dev)
        pix = torch.randint(0, npix, (4096,), generator=gen, device=dev)
        b64 = bt[:, pix].double()
        if args.mode == "gain":
            ph = g[cols, 0, :nd].double() @ b64
            ax = 10.0 ** (g[cols, 1, :nd].double() @ b64)
            ay = 10.0 ** (g[cols, 2, :nd].double() @ b64)
            want = torch.stack([ax * ph.cos(), ax * ph.sin(), ay * ph.cos(), ay * ph.sin()], 1)
            scale = torch.stack([ax, ax, ay, ay], 1)
        else:
            ph = g[cols, :nd].double() @ b64
            if args.mode == "tec":
                want, scale = ph[:, None], torch.ones_like(ph[:, None])
            else:
                want = torch.stack([ph.cos(), ph.sin(), ph.cos(), ph.sin()], 1)[:, :planes]
                scale = torch.ones_like(want)
        got = last[cols][:, :, pix].double()
        err = ((got - want).abs() / scale).max().item()
        line += f", max err {err:.2e}"
    print(line, flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""
Benchmark of the hot path h5parm solutions -> a-term cube (BASELINE.json metric: screen
voxels/s, voxel = pixel x antenna x time x freq = 16 bytes of cube).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload lofar|voronoi1k|skalow|small]

Workload at N=1 (default): BASELINE.json configs[2] -- synthetic LOFAR-HBA shape, 62 stations,
45 directions, 24 freqs, 512x512 grid.  One STEP = one pass of the hot path over a slab of
``--slots`` time slots (default 4 -> 5952 (time, freq, station) columns, 1.56e9 voxels,
25 GB of cube): KL = batched FP64 fit of the slab's systems + coefficient GEMV + the
basis-contraction/sincos kernel writing the cube slab; Voronoi = FP64 cos/sin LUT + fill
(and smoothing for the voronoi1k workload).  Time slabs are independent, so with N ranks each
rank owns its own time slots (weak scaling, no data-path collective; one tiny all-reduce of
per-antenna weight sums picks the reference station).

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(1, os.path.join(ROOT, "tests"))        # synthetic_cases: the seeded input generator

FIT_AHEAD = 2
FIT_GROUP = 4      # slabs per fit call of the streaming KL path
METRIC = "screen voxels/s (pix*ant*time*freq), KL & Voronoi"
UNIT = "voxel/s"
BYTES_PER_VOXEL = 16


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="lofar", choices=["lofar", "voronoi1k", "skalow", "small"])
    ap.add_argument("--slots", type=int, default=0, help="time slots per step (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--variant", type=int, default=-1, help="ssf_kl_eval variant (-1 auto)")
    ap.add_argument("--smooth-pix", type=float, default=-1.0, help="override sigma in pixels")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path, encoding="utf8") as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(kernel, workload="lofar"):
    """DRAM bytes per launch of `kernel` on `workload` from the committed ncu capture, if any."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        with open(path, encoding="utf8") as fh:
            return json.load(fh).get(f"{workload}:{kernel}")
    except (OSError, ValueError):
        return None


class ClockSampler:
    """SM clocks / throttle reasons sampled DURING the timed region: in-process NVML
    (nvidia_ml_py) every 50 ms from a helper thread; `nvidia-smi -lms` as the fallback.  One
    nvidia-smi process per rank on an 8-GPU box takes seconds to start and contends for the
    driver with the ranks' launches, NVML in-process does neither."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    # nvmlClocksEventReason* bit masks
    REASONS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20,
               "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.index = index
        self.rows = []          # (stamp, sm_mhz, sm_max_mhz, [reason names])
        self._proc = None
        self._thread = None
        self._stop = threading.Event()
        self._nvml = None

    def start(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[self.index]) if visible and visible.replace(",", "").isdigit() else self.index
            handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            smax = float(pynvml.nvmlDeviceGetMaxClockInfo(handle, pynvml.NVML_CLOCK_SM))
            self._nvml = (pynvml, handle, smax)
            self._thread = threading.Thread(target=self._poll_nvml, daemon=True)
            self._thread.start()
            return
        except Exception:  # noqa: BLE001 - any NVML problem -> nvidia-smi
            self._nvml = None
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self._thread = threading.Thread(target=self._pump, daemon=True)
        self._thread.start()

    def _poll_nvml(self):
        pynvml, handle, smax = self._nvml
        while not self._stop.is_set():
            try:
                sm = float(pynvml.nvmlDeviceGetClockInfo(handle, pynvml.NVML_CLOCK_SM))
                try:
                    bits = pynvml.nvmlDeviceGetCurrentClocksEventReasons(handle)
                except Exception:  # noqa: BLE001 - older binding name
                    bits = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(handle)
                self.rows.append((time.time(), sm, smax, [n for n, m in self.REASONS.items() if bits & m]))
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.05)

    def _pump(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self._proc.stdout:
            cells = [c.strip() for c in line.strip().split(",")]
            try:
                sm, smax = float(cells[0]), float(cells[1])
            except (ValueError, IndexError):
                continue
            why = [n for n, c in zip(names, cells[3:7]) if c.lower().startswith("active")]
            self.rows.append((time.time(), sm, smax, why))

    def stop(self, t_begin, t_end):
        if self._thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no clock source (NVML, nvidia-smi) available"]}
        if self._nvml is None:
            time.sleep(0.15)
            self._proc.terminate()
        self._stop.set()
        sm, smax, reasons = [], None, set()
        for stamp, mhz, mx, why in list(self.rows):
            if not t_begin - 0.05 <= stamp <= t_end + 0.15:
                continue
            sm.append(mhz)
            smax = mx
            reasons.update(why)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "samples": len(sm), "reasons": sorted(reasons),
                "source": "nvml" if self._nvml is not None else "nvidia-smi"}


_ORIG_AFFINITY = None


def bind_to_gpu_numa_node(index):
    """Pin this rank's threads to the CPUs NVML names as local to its GPU, before any pinned
    host buffer is allocated: first-touch then places the e2e staging buffers on the GPU's own
    NUMA node instead of sending every D2H byte across the socket interconnect.  Undone by
    ``restore_affinity`` before the CPU baseline (which wants all cores)."""
    global _ORIG_AFFINITY
    try:
        import pynvml

        if _ORIG_AFFINITY is None:
            _ORIG_AFFINITY = os.sched_getaffinity(0)
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(visible.split(",")[index]) if visible and visible.replace(",", "").isdigit() else index
        handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
        cpus &= _ORIG_AFFINITY
        if cpus:
            os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:  # noqa: BLE001 - affinity is an optimisation, never a requirement
        return None


def restore_affinity():
    if _ORIG_AFFINITY is not None:
        try:
            os.sched_setaffinity(0, _ORIG_AFFINITY)
        except OSError:
            pass


# ---------------------------------------------------------------------------- workload
def auto_slots(workload):
    return {"lofar": 4, "voronoi1k": 1, "skalow": 1, "small": 4}[workload]


def build_case(args, rank, total_slots):
    import synthetic_cases as synthetic

    kw = {}
    if args.workload == "skalow":
        kw["n_freqs"] = 1          # one (time, freq) slice of 512 stations = 34 GB of cube per step
    return synthetic.make_case(args.workload, seed=3, n_times=total_slots,
                               time_offset=rank * total_slots, **kw)


def make_screens(case, eng, variant, weight_sums):
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    kl = KLScreen("bench_kl", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    kl._engine = eng
    kl.eval_variant = variant
    kl.fit_ahead = FIT_AHEAD
    kl.fit_group = FIT_GROUP
    kl.setup(case.solutions, antenna_weight_sums=weight_sums)
    kl._prepare(case.cellsize_deg, None)
    vo = VoronoiScreen("bench_vo", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    vo._engine = eng
    vo.template_mode = "nearest"
    vo._has_nan = bool(np.isnan(np.asarray(case.solutions.val)).any())
    vo.patch_radec = case.patch_radec
    vo.source_names = case.solutions.dir
    vo.freqs_ph, vo.times_ph, vo.station_names = case.solutions.freq, case.solutions.time, case.solutions.ant
    vo._prepare(case.cellsize_deg, None)
    return kl, vo


def run_ours(args):
    import torch

    from ska_sdp_screen_fitting_b200 import _build
    from ska_sdp_screen_fitting_b200.engine import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    bind_to_gpu_numa_node(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not os.path.exists(_build.LIB_PATH):
        if rank == 0:
            _build.build()
        if dist is not None:
            dist.barrier()
    eng = Engine(device=f"cuda:{local}")
    slots = args.slots or auto_slots(args.workload)
    n_steps = args.warmup + args.steps
    # (FIT_AHEAD + 2) fit groups more than are evaluated: the pipelined KL path fits FIT_GROUP
    # slabs per call and keeps FIT_AHEAD groups (+ the one in flight) in front of the evaluation
    # through a bounded queue, so in steady state the fit never runs out of slabs during the K
    # timed steps: they contain K slab evaluations and (to within one group) K slab fits
    case = build_case(args, rank, slots * (n_steps + (FIT_AHEAD + 2) * FIT_GROUP))
    sol = case.solutions
    n_t, n_f, n_a, n_d = sol.val.shape
    n = case.n_pix
    cols = slots * n_f * n_a
    voxels_per_step = cols * n * n
    smooth_pix = args.smooth_pix if args.smooth_pix >= 0 else (5.0 if args.workload == "voronoi1k" else 0.0)

    # inputs resident in HBM before the timed region
    val_dev = eng.tensor(sol.val, torch.float64)
    wt_dev = eng.tensor(sol.weight, torch.float32)
    # metadata exchange: per-antenna weight sums -> reference station (global over time)
    sums = wt_dev.double().sum(dim=(0, 1, 3))
    if dist is not None:
        dist.all_reduce(sums)
    kl, vo = make_screens(case, eng, args.variant, sums.cpu().numpy())
    ring = [eng.empty((cols, 4, n, n), torch.float32) for _ in range(2)]

    kl_stream = kl.stream_slabs(val_dev, wt_dev, slots, ring, case.cellsize_deg)

    def kl_step(i):
        # the fits of the next group of slabs (side stream) overlap the evaluation of slab i
        k, _ = next(kl_stream)
        assert k == i

    def vo_step(i):
        t0 = i * slots
        v = val_dev[t0:t0 + slots]
        ph = (v - v[:, :, kl.ref_ind:kl.ref_ind + 1, :]).reshape(-1, n_d)
        vo._fill_columns(ph, ring[i % 2], case.cellsize_deg, None, smooth_pix, True)

    kept_timers = None

    def timed(step_fn, timer_keys):
        for i in range(args.warmup):
            step_fn(i)
        eng.check_fit_status()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        sampler = ClockSampler(local)
        sampler.start()
        time.sleep(0.25)
        eng.timers = {}
        eng.host_times = {}
        kl.timings.pop("fit_host_s", None)
        kl.timings.pop("fit_calls", None)
        kl.timings.pop("fit_slabs", None)
        launches0 = eng.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # the streamed KL fit runs a fixed number of slab groups ahead of the evaluation: bring
        # that lead to its steady-state value on both sides of the timed region (and order the
        # main stream after it), so the region holds as many slab fits as slab evaluations
        kl.settle_fit_lead()
        t_begin = time.time()
        ev0.record()
        for i in range(args.warmup, n_steps):
            step_fn(i)
        kl.settle_fit_lead()
        ev1.record()
        torch.cuda.synchronize()
        t_end = time.time()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        clocks = sampler.stop(t_begin, t_end)
        ms = ev0.elapsed_time(ev1)
        launches = eng.launch_count() - launches0
        kern = eng.timer_summary()
        nonlocal kept_timers
        kept_timers = eng.timers
        eng.timers = None
        eng.check_fit_status()
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=eng.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, {k: kern[k] for k in timer_keys if k in kern}, clocks

    peak, peak_src = measured_peaks()
    out = {}
    if args.workload != "voronoi1k":
        ms, launches, kern, clocks = timed(kl_step, ["kl_eval", "kl_patterns_update", "kl_flag_outliers", "kl_fit_pass"])
        # side-stream spans (queueing behind the evaluation's CTAs + execution), whole run incl. warm-up for the host part
        fit_breakdown = {k: {"calls": kern[k][0], "mean_ms": kern[k][1]} for k in kern if k != "kl_eval"}
        fit_breakdown["pattern_slots_in_use"] = kl._patterns.stats()[0]
        fit_breakdown["patterns_update_spans_ms"] = [round(a.elapsed_time(b), 3) for a, b in
                                                     (kept_timers or {}).get("kl_patterns_update", [])]
        fit_breakdown["host_s"] = dict(eng.host_times)
        fit_breakdown["torch_device_allocs_total"] = torch.cuda.memory_stats().get("num_device_alloc", None)
        kl_stream.close()                     # stops the fit helper thread
        torch.cuda.synchronize()
        n_launch, k_ms = kern["kl_eval"]
        achieved = voxels_per_step * BYTES_PER_VOXEL / (k_ms * 1e-3) / 1e9
        # the same kernel timed alone (no fit running beside it on the side stream): explains
        # how much of the in-step duration is sharing SMs with the overlapped FP64 fit
        white0 = kl.fit_device(val_dev[:slots], wt_dev[:slots], check=False)[0].view(-1, n_d)
        coeffs0 = eng.kl_coeffs(white0, kl._c_full)
        torch.cuda.synchronize()
        alone = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            eng.kl_eval(kl._basis, coeffs0, n_d, ring[0], args.variant)
            b.record()
            torch.cuda.synchronize()
            alone.append(a.elapsed_time(b))
        alone_ms = sorted(alone)[1]
        tc = args.variant != 0 and (n * n) % 128 == 0
        kname = (("contract_tc_bigk_kernel (ssf_kl_eval, tcgen05 split-TF32, persistent, basis operand in TMEM)"
                  if n_d > 56 else "contract_tc_kernel<KL> (ssf_kl_eval, tcgen05 split-TF32)") if tc
                 else "contract_tile_kernel<KL> (ssf_kl_eval, FP32 SIMT)")
        out["kl"] = dict(ms=ms, launches=launches, clocks=clocks,
                         fit_host_ms=kl.timings.get("fit_host_s", 0.0) / max(kl.timings.get("fit_slabs", 1), 1) * 1e3,
                         fit_slabs=kl.timings.get("fit_slabs", 0), fit_breakdown=fit_breakdown,
                         roofline={"bound": "hbm", "kernel": kname,
                                   "alone": {"kernel_ms": alone_ms,
                                             "achieved": voxels_per_step * BYTES_PER_VOXEL / (alone_ms * 1e-3) / 1e9,
                                             "frac": voxels_per_step * BYTES_PER_VOXEL / (alone_ms * 1e-3) / 1e9 / peak},
                                   "achieved": achieved, "peak": peak, "unit": "GB/s",
                                   "frac": achieved / peak, "traffic": recorded_traffic("kl_eval", args.workload),
                                   "kernel_ms": k_ms, "kernel_launches": n_launch, "peak_source": peak_src,
                                   "kernel_share_of_step": k_ms * n_launch / ms})
    ms_v, launches_v, kern_v, clocks_v = timed(vo_step, ["voronoi_fill", "gauss_smooth", "voronoi_contract"])
    vkey = "voronoi_contract" if "voronoi_contract" in kern_v else "voronoi_fill"
    n_launch, k_ms = kern_v[vkey]
    achieved_v = voxels_per_step * BYTES_PER_VOXEL / (k_ms * 1e-3) / 1e9
    out["voronoi"] = dict(ms=ms_v, launches=launches_v, clocks=clocks_v,
                          roofline={"bound": "hbm",
                                    "kernel": ("contract_tc_kernel<Voronoi> (ssf_voronoi_contract: smoothed cell "
                                               "indicators x cos/sin LUT, tcgen05 split-TF32)"
                                               if vkey == "voronoi_contract" else "fill_kernel (ssf_voronoi_fill)"),
                                    "achieved": achieved_v, "peak": peak, "unit": "GB/s",
                                    "frac": achieved_v / peak, "traffic": recorded_traffic(vkey, args.workload),
                                    "kernel_ms": k_ms, "kernel_launches": n_launch, "peak_source": peak_src,
                                    "kernel_share_of_step": k_ms * n_launch / ms_v})
    if "gauss_smooth" in kern_v:
        g_n, g_ms = kern_v["gauss_smooth"]
        # smoothing reads and writes the slab once per pass (2 passes): 4 x 16 B per voxel
        out["voronoi"]["smooth"] = {"sigma_pix": smooth_pix, "kernel_ms": g_ms,
                                    "achieved": voxels_per_step * 64 / (g_ms * 1e-3) / 1e9,
                                    "unit": "GB/s", "algorithmic_bytes_per_voxel": 64}
    if args.workload != "voronoi1k":
        # gain solutions (scalar phase + XX/YY amplitude screens, kl_screen.py:319-378): the
        # evaluation of three screens per voxel on the same kernel.  The amplitude fit runs once
        # per run on the slow amplitude grid, so the per-slab work is this evaluation next to the
        # phase fit measured above; the screens here are seeded log10 amplitudes of +-5 %.
        gen = torch.Generator(device=eng.device).manual_seed(7)
        white_g = kl.fit_device(val_dev[:slots], wt_dev[:slots], check=False)[0].view(-1, n_d)
        amp_fit = 0.02 * torch.randn((cols, n_d, 2), dtype=torch.float64, device=eng.device, generator=gen)
        amp_white = torch.einsum("de,cep->cdp", kl._pinv_full, amp_fit).contiguous()

        def gain_step(i):
            kl._fill_columns(white_g, ring[i % 2], case.cellsize_deg, None, 0.0, True, amp=amp_white)

        ms_g, launches_g, kern_g, clocks_g = timed(gain_step, ["kl_eval"])
        n_launch, k_ms = kern_g["kl_eval"]
        achieved_g = voxels_per_step * BYTES_PER_VOXEL / (k_ms * 1e-3) / 1e9
        out["gain"] = dict(ms=ms_g, launches=launches_g, clocks=clocks_g,
                           roofline={"bound": "hbm", "kernel": "contract_tc_kernel<gain> (ssf_kl_eval_gain: phase + "
                                     "log10 Axx + log10 Ayy screens, tcgen05 split-TF32)",
                                     "achieved": achieved_g, "peak": peak, "unit": "GB/s", "frac": achieved_g / peak,
                                     "traffic": recorded_traffic("kl_eval_gain", args.workload), "kernel_ms": k_ms,
                                     "kernel_launches": n_launch, "peak_source": peak_src,
                                     "kernel_share_of_step": k_ms * n_launch / ms_g})
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, eng, case, kl, vo, dist, smooth_pix)
    return dict(world=world, rank=rank, slots=slots, cols=cols, n=n, n_d=n_d, n_f=n_f, n_a=n_a,
                voxels_per_step=voxels_per_step, smooth_pix=smooth_pix, out=out, e2e=e2e, dist=dist,
                case=case)


def run_e2e(args, eng, case, kl, vo, dist, smooth_pix):
    """Same metric through the host-buffer boundary: every step copies that step's solutions
    (val, weight) from pinned host memory to the device, runs fit + evaluation, and copies the
    finished cube slab back into pinned host memory (where the FITS writer would take it).
    The D2H copy of step k runs on a second stream while step k+1 computes."""
    import torch

    sol = case.solutions
    n_t, n_f, n_a, n_d = sol.val.shape
    n = case.n_pix
    e_slots = 1
    e_steps = max(2, min(args.steps, 4))
    e_warm = 1
    cols = e_slots * n_f * n_a
    slab_elems = cols * 4 * n * n
    host_val = torch.from_numpy(np.ascontiguousarray(sol.val[:e_slots * (e_steps + e_warm)])).pin_memory()
    host_wt = torch.from_numpy(np.ascontiguousarray(sol.weight[:e_slots * (e_steps + e_warm)])).pin_memory()
    host_out = [torch.empty(slab_elems, dtype=torch.float32).pin_memory() for _ in range(2)]
    dev_out = [eng.empty((cols, 4, n, n), torch.float32) for _ in range(2)]
    copy_stream = torch.cuda.Stream()
    main = torch.cuda.current_stream()
    copied = [None, None]
    use_kl = args.workload != "voronoi1k"

    def step(i):
        b = i % 2
        t0 = i * e_slots
        v = host_val[t0:t0 + e_slots].to(eng.device, non_blocking=True)
        w = host_wt[t0:t0 + e_slots].to(eng.device, non_blocking=True)
        if copied[b] is not None:
            main.wait_event(copied[b])                 # slab b must have left the device
        if use_kl:
            white, _, _, _ = kl.fit_device(v, w, check=False)
            kl._fill_columns(white.view(-1, n_d), dev_out[b], case.cellsize_deg, None, 0.0, True)
        else:
            ph = (v - v[:, :, kl.ref_ind:kl.ref_ind + 1, :]).reshape(-1, n_d)
            vo._fill_columns(ph, dev_out[b], case.cellsize_deg, None, smooth_pix, True)
        done = torch.cuda.Event()
        done.record(main)
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(done)
            host_out[b].copy_(dev_out[b].view(-1), non_blocking=True)
            copied[b] = torch.cuda.Event()
            copied[b].record(copy_stream)

    for i in range(e_warm):
        step(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    t_begin = time.perf_counter()
    for i in range(e_warm, e_warm + e_steps):
        step(i)
    torch.cuda.synchronize()
    seconds = time.perf_counter() - t_begin
    eng.check_fit_status()
    if dist is not None:
        t = torch.tensor([seconds], dtype=torch.float64, device=eng.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        seconds = float(t.item())
    world = dist.get_world_size() if dist is not None else 1
    checksum = float(host_out[(e_warm + e_steps - 1) % 2][::4097].double().sum())
    return {"value": world * e_steps * cols * n * n / seconds, "unit": UNIT,
            "h2d_bytes_per_step": int(e_slots * n_f * n_a * n_d * 12),
            "d2h_bytes_per_step": int(slab_elems * 4), "steps": e_steps,
            "ms_per_step": seconds / e_steps * 1e3, "columns_per_step": cols,
            "path": "KLScreen.fit_device + ssf_kl_eval" if use_kl else "VoronoiScreen fill",
            "overlap": "D2H of step k on a copy stream under the compute of step k+1",
            "checksum": checksum}


# ---------------------------------------------------------------------------- CPU arm
def cpu_sample(case, use_kl, smooth_pix, n_slots, ncpu):
    """One bounded sample of the workload on the host cores, through the LOOP-FAITHFUL
    restatement of the reference (oracle/port/refloops.py): ONE (freq, station) pair x
    ``n_slots`` time slots at the full image size.  KL: fit of those systems
    (_process_station restated, serial like one worker of the reference) + the Pool-per-call
    per-pixel evaluation; Voronoi: serial fill + gaussian_filter + NaN fix.
    Returns (seconds, voxels)."""
    from oracle.port import klfit, refloops, screens, wcs as owcs

    sol = case.solutions
    n = case.n_pix
    f_ind, a_ind = 0, min(1, sol.val.shape[2] - 1)
    val = np.asarray(sol.val[:n_slots], np.float64)
    wt = np.asarray(sol.weight[:n_slots])
    t0 = time.perf_counter()
    if use_kl:
        ref = klfit.get_reference_station(wt, 10)
        a_fit = a_ind if a_ind != ref else a_ind + 1
        out = klfit.run_phase(val, wt, np.asarray(sol.source_positions), np.asarray(sol.station_positions),
                              ref, min(20, val.shape[3] - 1), stations=[a_fit], freqs=[f_ind])
        x, y = owcs.kl_grid_coords(case.rad, case.dec, case.width_deg, case.cellsize_deg,
                                   out["midra"], out["middec"])
        data = refloops.kl_make_matrix(x, y, out["pp"], out["screen"][:, f_ind, a_fit, :], ncpu=ncpu)
        cube = refloops.smooth_and_fix(data.astype(np.float32), 0.0)
    else:
        from oracle.port import template as otemplate
        px, py = owcs.voronoi_patch_pixels(case.patch_radec[:, 0], case.patch_radec[:, 1], case.rad,
                                           case.dec, n, case.cellsize_deg)
        tmpl = screens.nearest_patch_template(px, py, n)
        ph = screens.voronoi_reference_phases(val, 0)[:, f_ind, a_ind, :]
        data = refloops.voronoi_make_matrix(tmpl, ph)
        cube = refloops.smooth_and_fix(data.astype(np.float32), smooth_pix)
    seconds = time.perf_counter() - t0
    assert cube.shape == (n_slots, 4, n, n)
    return seconds, n_slots * n * n


def cpu_slots_for(use_kl, ncpu):
    # KL: one task per time slot, one task ~ n^2 x 14 us; give every core one slot.
    # Voronoi: serial; ~1e7 voxel/s -> 32 slots of 1024^2 in a few seconds.
    if use_kl:
        return max(1, min(ncpu, 256))
    return 32


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (loop-faithful
    oracle port; the Python reference cannot travel to the GPU box), all host threads it can
    use, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    import synthetic_cases as synthetic

    ncpu = os.cpu_count() or 1
    use_kl = args.workload != "voronoi1k"
    smooth_pix = args.smooth_pix if args.smooth_pix >= 0 else (5.0 if args.workload == "voronoi1k" else 0.0)
    n_slots = cpu_slots_for(use_kl, ncpu)
    case = synthetic.make_case(args.workload, seed=3, n_times=n_slots, n_freqs=1,
                               n_stations=min(4, synthetic.SHAPES[args.workload][0]))
    for _ in range(min(args.warmup, 1)):
        cpu_sample(case, use_kl, smooth_pix, max(1, n_slots // 8), ncpu)
    total_s, total_v = 0.0, 0
    for _ in range(args.steps):
        s, v = cpu_sample(case, use_kl, smooth_pix, n_slots, ncpu)
        total_s += s
        total_v += v
    value = total_v / total_s
    sample = (f"1 (freq, station) pair x {n_slots} time slots at {case.n_pix}x{case.n_pix}, "
              f"{len(case.solutions.dir)} directions, per step; "
              + ("fit (serial) + Pool(ncpu) per-pixel KL evaluation" if use_kl else
                 f"serial fill + gaussian_filter sigma={smooth_pix} px + NaN fix"))
    return {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference",
            "config": workload_config(args, case.n_pix, len(case.solutions.dir), None, None),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": ncpu if use_kl else 1,
                             "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def workload_config(args, n, n_d, cols, slots):
    names = {"lofar": "synthetic LOFAR-HBA shape: 62 stations, 45 directions, 24 freqs, 512x512 grid, KL + Voronoi",
             "voronoi1k": "synthetic Voronoi with Gaussian smoothing at 1024x1024 grid, 62 stations, 45 directions, 24 freqs",
             "skalow": "synthetic SKA-Low shape: 512 stations, 100 directions, 2048x2048 grid, KL (one freq per step)",
             "small": "small synthetic test shape"}
    cfg = {"workload": names[args.workload], "image": f"{n}x{n}", "directions": n_d}
    if cols is not None:
        cfg.update({"time_slots_per_step": slots, "columns_per_step": cols,
                    "cache": f"each step writes {cols * n * n * 16 / 1e9:.1f} GB of cube into a 2-slab "
                             "ring (far larger than the 126 MB L2), no explicit flush needed",
                    "sharding": "time slabs per rank, no data-path collective"})
    return cfg


def main():
    args = parse_args()
    if args.impl == "reference":
        line = run_reference(args)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    res = run_ours(args)
    dist = res["dist"]
    if res["rank"] == 0:
        out = res["out"]
        head = "kl" if "kl" in out else "voronoi"
        ms = out[head]["ms"]
        world = res["world"]
        value = world * args.steps * res["voxels_per_step"] / (ms * 1e-3)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "path": "KL (fit + evaluation)" if head == "kl" else "Voronoi (fill + smoothing)",
                "config": workload_config(args, res["n"], res["n_d"], res["cols"], res["slots"]),
                "roofline": out[head]["roofline"], "gpu_launches": out[head]["launches"],
                "clocks": out[head]["clocks"]}
        if "fit_host_ms" in out[head]:
            line["fit_host_ms_per_slab"] = out[head]["fit_host_ms"]
            line["fit_slabs_in_timed_region"] = out[head]["fit_slabs"]
            line["fit_breakdown"] = out[head]["fit_breakdown"]
        if head == "kl":
            vms = out["voronoi"]["ms"]
            line["voronoi"] = {"value": world * args.steps * res["voxels_per_step"] / (vms * 1e-3),
                               "unit": UNIT, "ms_per_step": vms / args.steps,
                               "smooth_pix": res["smooth_pix"], "roofline": out["voronoi"]["roofline"],
                               "gpu_launches": out["voronoi"]["launches"], "clocks": out["voronoi"]["clocks"]}
        elif "smooth" in out["voronoi"]:
            line["smooth"] = out["voronoi"]["smooth"]
        if "gain" in out:
            gms = out["gain"]["ms"]
            line["kl_gain_eval"] = {"value": world * args.steps * res["voxels_per_step"] / (gms * 1e-3),
                                    "unit": UNIT, "ms_per_step": gms / args.steps,
                                    "roofline": out["gain"]["roofline"], "gpu_launches": out["gain"]["launches"]}
        if res["e2e"] is not None:
            line["e2e"] = res["e2e"]
        if world == 1 and not args.no_cpu_baseline:
            restore_affinity()
            ncpu = os.cpu_count() or 1
            use_kl = head == "kl"
            import synthetic_cases as synthetic
            n_slots = cpu_slots_for(use_kl, ncpu)
            case = synthetic.make_case(args.workload, seed=3, n_times=n_slots, n_freqs=1,
                                       n_stations=min(4, synthetic.SHAPES[args.workload][0]))
            secs, vox = cpu_sample(case, use_kl, res["smooth_pix"], n_slots, ncpu)
            line["cpu_baseline"] = {
                "value": vox / secs, "unit": UNIT, "cores": ncpu if use_kl else 1, "kind": "port",
                "seconds": secs,
                "sample": f"1 (freq, station) pair x {n_slots} time slots at {res['n']}x{res['n']}, "
                          f"{res['n_d']} directions: loop-faithful port of the reference "
                          "(oracle/port/refloops.py), " + ("serial fit + Pool(ncpu) per-pixel evaluation"
                                                           if use_kl else "serial fill + gaussian_filter")}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

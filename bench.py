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
import gc
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

# Fit groups the streamed fit keeps in front of the evaluation.  6 groups = 24 LOFAR steps = ~100 ms of
# slack: the thread that issues the fits is descheduled for 20-80 ms now and then on these (virtual)
# hosts -- wall time of one stage 70 ms with 0.5 ms of CPU time, with or without the NVML poller,
# the collector, allocations -- and with a lead of 2 groups (35 ms) that stalled the GPU: one
# bench run in ~10 read 6-7 ms per step instead of 4.3.
FIT_AHEAD = int(os.environ.get("SSF_BENCH_FIT_AHEAD", "6"))
FIT_GROUP = 4      # slabs per fit call of the streaming KL path
# Slabs over which the SKA-Low KL pass warms the flag-pattern cache before the timed steps.  New
# patterns per 512-system slab at 1 % flags (simulated): 138 cold, 76 after 48 slabs, 53 after 100,
# 42 after 200, 35-40 from 300 on; a 65 536-slot cache fills (and starts over) every ~1500 slabs, so
# the average over a 65 536-slab job is ~44 per slab -- what the timed steps see after 192 slabs.
KL_WARM_SKALOW = int(os.environ.get("SSF_BENCH_KL_WARM", "192"))
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
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the voronoi1k / skalow sub-runs appended to the default lofar line")
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
    (nvidia_ml_py) every 20 ms from a helper thread; `nvidia-smi -lms` as the fallback.  One
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
        period = float(os.environ.get("SSF_BENCH_NVML_PERIOD", "0.02"))
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
            self._stop.wait(period)

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


def build_case(workload, rank, total_slots):
    import synthetic_cases as synthetic

    kw = {}
    if workload == "skalow":
        kw["n_freqs"] = 1          # one (time, freq) slice of 512 stations = 34 GB of cube per step
    return synthetic.make_case(workload, seed=3, n_times=total_slots, time_offset=rank * total_slots, **kw)


def make_screens(case, eng, variant, weight_sums):
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    setup = {}
    t0 = time.perf_counter()
    kl = KLScreen("bench_kl", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    kl._engine = eng
    kl.eval_variant = variant
    kl.fit_ahead = FIT_AHEAD
    kl.fit_group = FIT_GROUP
    kl.setup(case.solutions, antenna_weight_sums=weight_sums)
    _sync()
    setup["kl_setup_s (reference station, orders, pierce points, full-pattern decomposition = pattern slot 0)"] = \
        time.perf_counter() - t0
    t0 = time.perf_counter()
    kl._prepare(case.cellsize_deg, None)
    _sync()
    setup["kl_basis_s (ssf_kl_basis, FP64, once per run)"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    vo = VoronoiScreen("bench_vo", None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    vo._engine = eng
    vo.template_mode = "nearest"
    vo._has_nan = bool(np.isnan(np.asarray(case.solutions.val)).any())
    vo.patch_radec = case.patch_radec
    vo.source_names = case.solutions.dir
    vo.freqs_ph, vo.times_ph, vo.station_names = case.solutions.freq, case.solutions.time, case.solutions.ant
    vo._prepare(case.cellsize_deg, None)
    _sync()
    setup["voronoi_template_s (ssf_voronoi_assign)"] = time.perf_counter() - t0
    return kl, vo, setup


def _sync():
    import torch

    torch.cuda.synchronize()


def measure_workload(args, eng, dist, workload, steps, warmup, rank, local, parts):
    """Device-timed passes of the hot path over one workload, inputs resident in HBM.  ``parts``:
    which of "kl", "voronoi", "gain", "kl_pairs" to run.  Returns a dict of result blocks."""
    import torch

    slots = args.slots or auto_slots(workload)
    # SKA-Low shape: at 1 % flags two thirds of the 100-direction systems carry a flag pattern and
    # a cold pattern cache decomposes ~490 matrices per slab; after some tens of slabs the one- and
    # two-flag patterns are cached and ~90 per slab are new.  A 2048-slot x 32-freq job has 65 536
    # slabs, so the steady state is what the timed region should see: the KL pass warms the cache
    # over KL_WARM_SKALOW slabs first (each warm-up step is a full fit + evaluation step).
    kl_warm = max(warmup, KL_WARM_SKALOW) if workload == "skalow" else warmup
    n_steps = kl_warm + steps
    # (FIT_AHEAD + 2) fit groups more than are evaluated: the pipelined KL path fits FIT_GROUP
    # slabs per call and keeps FIT_AHEAD groups (+ the one in flight) in front of the evaluation
    # through a bounded queue, so in steady state the fit never runs out of slabs during the K
    # timed steps: they contain K slab evaluations and (to within one group) K slab fits
    case = build_case(workload, rank, slots * (n_steps + (FIT_AHEAD + 2) * FIT_GROUP))
    sol = case.solutions
    n_t, n_f, n_a, n_d = sol.val.shape
    n = case.n_pix
    cols = slots * n_f * n_a
    voxels_per_step = cols * n * n
    smooth_pix = args.smooth_pix if args.smooth_pix >= 0 else (5.0 if workload == "voronoi1k" else 0.0)

    # inputs resident in HBM before the timed region
    val_dev = eng.tensor(sol.val, torch.float64)
    wt_dev = eng.tensor(sol.weight, torch.float32)
    # metadata exchange: per-antenna weight sums -> reference station (global over time)
    sums = wt_dev.double().sum(dim=(0, 1, 3))
    if dist is not None:
        dist.all_reduce(sums)
    kl, vo, setup = make_screens(case, eng, args.variant, sums.cpu().numpy())
    ring = [eng.empty((cols, 4, n, n), torch.float32) for _ in range(2)]
    peak, peak_src = measured_peaks()
    kept = {}

    def timed(step_fn, timer_keys, warmup=warmup):
        n_steps = warmup + steps
        for i in range(warmup):
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
        for key in ("fit_host_s", "fit_calls", "fit_rows"):
            kl.timings.pop(key, None)
        launches0 = eng.launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        # the streamed KL fit runs a fixed number of slab groups ahead of the evaluation: bring
        # that lead to its steady-state value on both sides of the timed region (and order the
        # main stream after it), so the region holds as many slab fits as slab evaluations
        kl.settle_fit_lead()
        # no collector pauses inside the timed steps (as timeit does): a generation-2 pass holds
        # the GIL for tens of ms and stalls the thread that issues the fits
        gc.collect()
        gc.disable()
        t_begin = time.time()
        ev0.record()
        try:
            for i in range(warmup, n_steps):
                step_fn(i)
            kl.settle_fit_lead()
            ev1.record()
            torch.cuda.synchronize()
        finally:
            gc.enable()
        t_end = time.time()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        clocks = sampler.stop(t_begin, t_end)
        ms = ev0.elapsed_time(ev1)
        launches = eng.launch_count() - launches0
        kern = eng.timer_summary()
        kept["timers"] = eng.timers
        eng.timers = None
        eng.check_fit_status()
        if dist is not None:
            t = torch.tensor([ms], dtype=torch.float64, device=eng.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches, {k: kern[k] for k in timer_keys if k in kern}, clocks

    def roof(kernel, k_ms, n_launch, ms, bytes_per_voxel=BYTES_PER_VOXEL, traffic_key=None):
        achieved = voxels_per_step * bytes_per_voxel / (k_ms * 1e-3) / 1e9
        line = {"bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": recorded_traffic(traffic_key, workload) if traffic_key else None,
                "algorithmic_bytes_per_voxel": bytes_per_voxel, "kernel_ms": k_ms, "kernel_launches": n_launch,
                "peak_source": peak_src, "kernel_share_of_step": k_ms * n_launch / ms}
        if achieved > peak:
            # the peak is a COPY bandwidth (half reads, half writes); these kernels only write, and a
            # write-only stream is not held to it: scripts/stream_write_probe.cu measures 7.0-7.5 TB/s
            # for stores alone on this part (profiles/r02_stream_write_probe.txt)
            line["note"] = "write-only kernel: above the measured COPY bandwidth, below the 7.5 TB/s of a bare store stream"
        return line

    tc = args.variant != 0 and (n * n) % 128 == 0
    big = n_d > 56
    out = {"setup": setup}

    def kl_run(out_flags, planes):
        slabs = ring if planes == 4 else [r.view(-1)[:cols * planes * n * n].view(cols, planes, n, n) for r in ring]
        stream = kl.stream_slabs(val_dev, wt_dev, slots, slabs, case.cellsize_deg, out_flags=out_flags)

        def kl_step(i):
            # the fits of the next group of slabs (side stream) overlap the evaluation of slab i
            k, _ = next(stream)
            assert k == i

        res = timed(kl_step, ["kl_eval", "kl_patterns_update", "kl_flag_outliers", "kl_fit_pass"], warmup=kl_warm)
        stream.close()                        # stops the fit helper thread
        torch.cuda.synchronize()
        return res

    if "kl" in parts:
        ms, launches, kern, clocks = kl_run(0, 4)
        # side-stream spans (queueing behind the evaluation's CTAs + execution)
        fit_breakdown = {k: {"calls": kern[k][0], "mean_ms": kern[k][1]} for k in kern if k != "kl_eval"}
        fit_breakdown["pattern_slots_in_use"] = kl._patterns.stats()[0]
        fit_breakdown["warmup_steps"] = kl_warm
        fit_breakdown["host_s"] = dict(eng.host_times)
        n_launch, k_ms = kern["kl_eval"]
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
        # the fit alone: one call over a group of FIT_GROUP slabs, as the streamed path issues it
        g_slots = slots * FIT_GROUP
        fit_alone = []
        for _ in range(3):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record()
            kl.fit_device(val_dev[:g_slots], wt_dev[:g_slots], check=False)
            b.record()
            torch.cuda.synchronize()
            fit_alone.append(a.elapsed_time(b))
        eng.check_fit_status()
        fit_ms = sorted(fit_alone)[1]
        kname = (("contract_tc_narrow_kernel (ssf_kl_eval, tcgen05 split-TF32, two CTAs per SM, A_hi in TMEM / A_lo in "
                  "shared memory, 32-column tiles)"
                  if big else "contract_tc_kernel<KL> (ssf_kl_eval, tcgen05 split-TF32)") if tc
                 else "contract_tile_kernel<KL> (ssf_kl_eval, FP32 SIMT)")
        r = roof(kname, k_ms, n_launch, ms, traffic_key="kl_eval")
        r["alone"] = {"kernel_ms": alone_ms,
                      "achieved": voxels_per_step * BYTES_PER_VOXEL / (alone_ms * 1e-3) / 1e9,
                      "frac": voxels_per_step * BYTES_PER_VOXEL / (alone_ms * 1e-3) / 1e9 / peak}
        out["kl"] = dict(ms=ms, launches=launches, clocks=clocks, roofline=r,
                         fit_host_ms=kl.timings.get("fit_host_s", 0.0) / max(kl.timings.get("fit_rows", 1) / slots, 1) * 1e3,
                         fit_slabs=kl.timings.get("fit_rows", 0) // slots, fit_breakdown=fit_breakdown,
                         fit={"systems_per_s_alone": g_slots * n_f * n_a / (fit_ms * 1e-3),
                              "systems_per_call": g_slots * n_f * n_a, "ms_per_call_alone": fit_ms,
                              "systems_per_s_in_step": cols / (ms / steps * 1e-3),
                              "note": "FP64, latency-bound; inside the step it runs on a side stream under the "
                                      "evaluation, so its in-step rate is the step rate"})
    if "kl_pairs" in parts:
        # the format the API path (Screen.write) asks the same kernel for when the solutions are
        # phase-only: planes 0, 1 only, stored big-endian -- 8 B per voxel leave the SM
        ms, launches, kern, clocks = kl_run(3, 2)
        n_launch, k_ms = kern["kl_eval"]
        out["kl_pairs"] = dict(ms=ms, launches=launches, clocks=clocks,
                               roofline=roof("same kernel, out_flags = SSF_OUT_BIG_ENDIAN | SSF_OUT_PLANES2 "
                                             "(plane pairs in FITS byte order)", k_ms, n_launch, ms, bytes_per_voxel=8))
    if "voronoi" in parts:
        def vo_step(i):
            t0 = i * slots
            v = val_dev[t0:t0 + slots]
            ph = (v - v[:, :, kl.ref_ind:kl.ref_ind + 1, :]).reshape(-1, n_d)
            vo._fill_columns(ph, ring[i % 2], case.cellsize_deg, None, smooth_pix, True)

        ms_v, launches_v, kern_v, clocks_v = timed(vo_step, ["voronoi_fill", "gauss_smooth", "voronoi_contract"])
        vkey = "voronoi_contract" if "voronoi_contract" in kern_v else "voronoi_fill"
        n_launch, k_ms = kern_v[vkey]
        out["voronoi"] = dict(ms=ms_v, launches=launches_v, clocks=clocks_v, smooth_pix=smooth_pix,
                              roofline=roof("contract_tc_kernel<Voronoi> (ssf_voronoi_contract: smoothed cell "
                                            "indicators x cos/sin LUT, tcgen05 split-TF32)"
                                            if vkey == "voronoi_contract" else "fill_kernel (ssf_voronoi_fill)",
                                            k_ms, n_launch, ms_v, traffic_key=vkey))
    if "voronoi_direct" in parts:
        # the reference's own order of operations (fill -> gaussian_filter -> NaN fix), the path of
        # NaN-carrying and gain slabs: 16 B written by the fill + 2 x (16 B read + 16 B written)
        # by the separable filter + the NaN-fix pass
        vo.smooth_mode = "direct"

        def vd_step(i):
            t0 = i * slots
            v = val_dev[t0:t0 + slots]
            ph = (v - v[:, :, kl.ref_ind:kl.ref_ind + 1, :]).reshape(-1, n_d)
            vo._fill_columns(ph, ring[i % 2], case.cellsize_deg, None, smooth_pix or 5.0, True)

        ms_d, launches_d, kern_d, clocks_d = timed(vd_step, ["voronoi_fill", "gauss_smooth"])
        vo.smooth_mode = "auto"
        g_n, g_ms = kern_d["gauss_smooth"]
        out["voronoi_direct"] = dict(ms=ms_d, launches=launches_d, clocks=clocks_d, smooth_pix=smooth_pix or 5.0,
                                     roofline=roof("ypass_kernel + xpass kernel (ssf_gauss_smooth, in place, "
                                                   "2 passes x (read + write))", g_ms, g_n, ms_d, bytes_per_voxel=64))
    if "gain" in parts:
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
        out["gain"] = dict(ms=ms_g, launches=launches_g, clocks=clocks_g,
                           roofline=roof("contract_tc_kernel<gain> (ssf_kl_eval_gain: phase + log10 Axx + log10 Ayy "
                                         "screens, tcgen05 split-TF32)", k_ms, n_launch, ms_g, traffic_key="kl_eval_gain"))
    del ring, val_dev, wt_dev
    torch.cuda.empty_cache()
    return dict(slots=slots, cols=cols, n=n, n_d=n_d, n_f=n_f, n_a=n_a, voxels_per_step=voxels_per_step,
                smooth_pix=smooth_pix, out=out, case=case, kl=kl, vo=vo)


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
    parts = {"lofar": ["kl", "kl_pairs", "voronoi", "gain"], "small": ["kl", "kl_pairs", "voronoi", "gain"],
             "voronoi1k": ["voronoi", "voronoi_direct"], "skalow": ["kl", "voronoi", "gain"]}[args.workload]
    res = measure_workload(args, eng, dist, args.workload, args.steps, args.warmup, rank, local, parts)
    res.update(world=world, rank=rank, dist=dist, extra={}, e2e=None)
    if not args.no_e2e:
        res["e2e"] = run_e2e(args, eng, res, dist, rank, world)
    res["kl"] = res["vo"] = None
    torch.cuda.empty_cache()
    if args.workload == "lofar" and not args.no_extras:
        # BASELINE.json configs[3] and configs[4] in the same driver-observed line: short sub-runs
        # with their own rooflines and clock samples
        import copy
        for name, sub_parts in (("voronoi1k", ["voronoi", "voronoi_direct"]), ("skalow", ["kl", "gain"])):
            sub_args = copy.copy(args)
            sub_args.slots, sub_args.smooth_pix = 0, -1.0
            steps, warm = min(args.steps, 5), 3
            sub = measure_workload(sub_args, eng, dist, name, steps, warm, rank, local, sub_parts)
            sub["kl"] = sub["vo"] = None
            sub.update(steps=steps, warmup=warm)
            res["extra"][name] = sub
            torch.cuda.empty_cache()
    return res


def tmpfs_dir(need_bytes):
    """A directory for the e2e output cube: tmpfs (/dev/shm) when it has room, else the system
    temp directory, else None (the run then stops at the pinned host buffer and says so)."""
    import shutil
    import tempfile

    for base in ("/dev/shm", tempfile.gettempdir()):
        try:
            free = shutil.disk_usage(base).free
        except OSError:
            continue
        if free > need_bytes * 1.3 + (4 << 30):
            return tempfile.mkdtemp(prefix="ssf_e2e_", dir=base), base, free
    return None, None, 0


def run_e2e(args, eng, res, dist, rank, world):
    """Same metric through the public API with HOST buffers: ``Screen.process()`` +
    ``Screen.write()`` (what ``make_aterm_image`` runs, make_aterm_images.py:146-153) on an
    in-memory ``Solutions`` record -- per block of time slots the H2D copy of its val / weight, the
    fit, the evaluation in FITS byte order, the D2H copy of the slab on a copy stream, and the
    placement of every byte in the output FITS file on tmpfs by threaded copies into a mapping of the file.  With N
    ranks the ranks share one cube and each writes the byte range of its own time slots."""
    import shutil

    import torch

    import synthetic_cases as synthetic
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    use_kl = args.workload != "voronoi1k"
    case0 = res["case"]
    n, n_d = res["n"], res["n_d"]
    n_f, n_a = res["n_f"], res["n_a"]
    per_slot = n_f * n_a * n * n                      # voxels per time slot
    e_warm = 1
    e_steps = max(2, min(args.steps, 4))              # time slots per rank in the timed write
    out_dir, base, free = (None, None, 0)
    if rank == 0:
        while e_steps > 1 and tmpfs_dir(world * e_steps * per_slot * 16)[0] is None:
            e_steps -= 1
        out_dir, base, free = tmpfs_dir(world * e_steps * per_slot * 16)
    if dist is not None:
        box = [out_dir, base, e_steps]
        dist.broadcast_object_list(box, src=0)
        out_dir, base, e_steps = box
    kw = {"n_freqs": 1} if args.workload == "skalow" else {}
    smooth_pix = res["smooth_pix"]

    def screen_for(n_slots, offset, name):
        case = synthetic.make_case(args.workload, seed=3, n_times=n_slots, time_offset=offset, **kw)
        cls = KLScreen if use_kl else VoronoiScreen
        scr = cls(name, None, None, case.rad, case.dec, case.width_deg, case.width_deg)
        scr._engine = eng
        scr.solutions = case.solutions                # host numpy arrays: the h5parm's val / weight
        if use_kl:
            scr.eval_variant = args.variant
        else:
            scr.patch_radec = case.patch_radec
            scr.template_mode = "nearest"
        if out_dir is None:
            scr.sink = "null"
        return scr, case

    target = out_dir or "/tmp"
    timings = {}
    # warm-up: a whole process() + write() of one slot per rank (loads kernels, pins the staging
    # buffers, builds the basis / template: all of it reported as `setup`, none of it timed below)
    scr, case = screen_for(world * e_warm, 0, "e2e_warm")
    t0 = time.perf_counter()
    scr.process()
    _sync()
    timings["process_s"] = time.perf_counter() - t0
    t0 = time.perf_counter()
    scr.write(target, case.cellsize_deg, smooth_pix=smooth_pix)
    timings["first_write_s"] = time.perf_counter() - t0
    timings["prepare_s"] = scr.timings.get("prepare_s")
    rings = scr._rings

    def one_run(sink):
        s2, c2 = screen_for(world * e_steps, world * e_warm, "e2e")
        s2._rings = rings                              # pinned staging buffers are per process, not per cube
        if use_kl:
            s2._basis, s2._basis_key = scr._basis, scr._basis_key
        else:
            s2._tmpl_dev, s2._tmpl_key, s2.data_rasertize_template = scr._tmpl_dev, scr._tmpl_key, scr.data_rasertize_template
            s2._ind_dev, s2._ind_key = scr._ind_dev, scr._ind_key
        if sink == "null":
            s2.sink = "null"
        _sync()
        if dist is not None:
            dist.barrier()
        t_begin = time.perf_counter()
        s2.process()
        t_proc = time.perf_counter() - t_begin
        s2.write(target, c2.cellsize_deg, smooth_pix=smooth_pix)
        _sync()
        seconds = time.perf_counter() - t_begin
        if dist is not None:
            t = torch.tensor([seconds], dtype=torch.float64, device=eng.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            seconds = float(t.item())
        return s2, seconds, t_proc

    s_null, sec_null, _ = one_run("null")
    s_file, sec_file, t_proc = one_run("file") if out_dir is not None else (s_null, sec_null, 0.0)
    tm = s_file.timings
    voxels = world * e_steps * per_slot
    checksum = None
    if out_dir is not None and rank == 0:
        from ska_sdp_screen_fitting_b200 import fits
        _, data = fits.read(os.path.join(out_dir, "e2e_0.fits"))
        checksum = float(np.asarray(data[-1, -1, -1, :, ::37, ::41], np.float64).sum())
        assert data.shape[0] == world * e_steps
    if dist is not None:
        dist.barrier()
    if out_dir is not None and rank == 0:
        shutil.rmtree(out_dir, ignore_errors=True)
    slab_cols = tm["columns_per_slab"]
    return {"value": voxels / sec_file, "unit": UNIT,
            "h2d_bytes_per_step": int(n_f * n_a * n_d * 12),
            "d2h_bytes_per_step": int(tm["d2h_bytes"] // e_steps), "steps": e_steps,
            "step": "one time slot per rank",
            "ms_per_step": sec_file / e_steps * 1e3, "columns_per_step": n_f * n_a,
            "path": ("KLScreen" if use_kl else "VoronoiScreen") + ".process() + Screen.write() on in-memory host "
                    "Solutions (the calls make_aterm_image makes); timed region = both calls",
            "sink": f"FITS cube on {base}" if out_dir is not None else "none: no file system with room, stops at the pinned host buffer",
            "file_bytes_per_step": int(tm["file_bytes"] // e_steps),
            "planes_copied": tm["planes_copied"], "out_flags": tm["out_flags"],
            "pipeline": {"columns_per_slab": slab_cols, "host_buffers": tm["host_buffers"],
                         "write_threads": tm["write_threads"], "process_s": t_proc,
                         "write_s": tm["write_s"], "placement_busy_s": tm["pwrite_s"],
                         "file_GBps_while_placing": tm["file_bytes"] / max(tm["pwrite_s"], 1e-9) / 1e9,
                         "wait_for_staging_buffer_s": tm["flush_wait_s"],
                         "placement": "ssf_host_place: threaded copies into a shared mapping of the FITS file, each "
                                      "plane pair placed twice",
                         "sink_limit": "first touch of tmpfs pages: 8-10 GB/s on these boxes whatever the method "
                                       "(profiles/r02_file_sink.txt); `without_file` is the same pipeline up to the "
                                       "pinned host buffer"},
            "without_file": {"value": voxels / sec_null, "seconds": sec_null,
                             "d2h_GBps": s_null.timings["d2h_bytes"] / sec_null / 1e9,
                             "note": "same pipeline, sink = pinned host buffer: separates the link from the file system"},
            "setup": {"process_s (first call: read rows, weight sums, pattern slot 0; kernels load)": timings["process_s"],
                      "prepare_s (basis / template, once per run)": timings["prepare_s"],
                      "first_write_s (incl. prepare and pinning the staging buffers)": timings["first_write_s"]},
            "overlap": "fit of block k+1 (side stream) | evaluation of slab k (main) | D2H of slab k-1 (copy stream) | "
                       "file placement of slab k-2 (flusher thread + write pool)",
            "checksum": checksum}


# ---------------------------------------------------------------------------- CPU arm
def reference_available():
    try:
        from oracle import refshim
        return refshim.available()
    except Exception:  # noqa: BLE001
        return False


def cpu_sample_reference(workload, use_kl, smooth_pix, n_slots, ncpu):
    """One bounded sample of the workload through the REFERENCE's own public API
    (``make_aterm_image`` of the unmodified reference modules under oracle.refshim): reference
    station + ONE fitted station, one frequency, ``n_slots`` time slots at the full image size.
    KL: stationscreen.run (serial over its one frequency) + Pool(ncpu) per-pixel evaluation per
    (freq, station) + NaN fix + FITS write; Voronoi: template + serial fill + gaussian_filter +
    NaN fix + FITS write.  Returns (seconds, voxels)."""
    import contextlib
    import io
    import shutil
    import tempfile

    import synthetic_cases as synthetic
    from oracle import refrun
    from oracle.refshim import fake_h5parm

    case = synthetic.make_case(workload, seed=3, n_times=n_slots, n_freqs=1, n_stations=2)
    tmp = tempfile.mkdtemp(prefix="ssf_ref_")
    try:
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            refrun.run_case(case, "kl" if use_kl else "tessellated", tmp,
                            smooth_deg=smooth_pix * case.cellsize_deg, ncpu=ncpu)
        seconds = time.perf_counter() - t0
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
        fake_h5parm.reset()
    return seconds, n_slots * 2 * case.n_pix * case.n_pix


def cpu_sample(case, use_kl, smooth_pix, n_slots, ncpu):
    """FALLBACK when no copy of the reference is available: one bounded sample through the
    LOOP-FAITHFUL restatement of the reference (oracle/port/refloops.py): ONE (freq, station)
    pair x ``n_slots`` time slots at the full image size.  Returns (seconds, voxels)."""
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
    # KL: one Pool task per time slot and (freq, station), one task ~ n^2 x 14 us; two stations are
    # evaluated one after the other, so give every core one slot.
    # Voronoi: serial; ~1e7 voxel/s -> 16 slots x 2 stations of 1024^2 in a few seconds.
    if use_kl:
        return max(1, min(ncpu, 256))
    return 16


def cpu_arm(workload, smooth_pix, ncpu):
    """(value, seconds, kind, cores, sample text) of the CPU baseline for ``workload``."""
    import synthetic_cases as synthetic

    use_kl = workload != "voronoi1k"
    n_slots = cpu_slots_for(use_kl, ncpu)
    st, nd, _, _, npx, _ = synthetic.SHAPES[workload]
    if reference_available():
        secs, vox = cpu_sample_reference(workload, use_kl, smooth_pix, n_slots, ncpu)
        kind = "reference"
        sample = (f"the reference's make_aterm_image (unmodified modules, oracle/_ref under oracle.refshim) on "
                  f"2 stations (reference + 1 fitted) x 1 freq x {n_slots} time slots at {npx}x{npx}, {nd} "
                  "directions: " + ("stationscreen.run + Pool(ncpu) per-pixel KL evaluation + NaN fix + FITS write"
                                    if use_kl else
                                    f"template + serial fill + gaussian_filter sigma={smooth_pix} px + NaN fix + FITS write"))
    else:
        case = synthetic.make_case(workload, seed=3, n_times=n_slots, n_freqs=1, n_stations=min(4, st))
        secs, vox = cpu_sample(case, use_kl, smooth_pix, n_slots, ncpu)
        kind = "port"
        sample = (f"1 (freq, station) pair x {n_slots} time slots at {npx}x{npx}, {nd} directions: loop-faithful "
                  "port of the reference (oracle/port/refloops.py; no copy of the reference on this box), "
                  + ("serial fit + Pool(ncpu) per-pixel evaluation" if use_kl else "serial fill + gaussian_filter"))
    return vox / secs, secs, kind, (ncpu if use_kl else 1), sample


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the box's host
    cores (the unmodified reference modules from oracle/_ref; the loop-faithful port only if no
    copy exists), all host threads it can use, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    import synthetic_cases as synthetic

    ncpu = os.cpu_count() or 1
    use_kl = args.workload != "voronoi1k"
    smooth_pix = args.smooth_pix if args.smooth_pix >= 0 else (5.0 if args.workload == "voronoi1k" else 0.0)
    total_s, total_v = 0.0, 0
    kind = cores = sample = None
    for i in range(min(args.warmup, 1) + args.steps):
        value, secs, kind, cores, sample = cpu_arm(args.workload, smooth_pix, ncpu)
        if i >= min(args.warmup, 1):
            total_s += secs
            total_v += value * secs
    value = total_v / total_s
    # the same `config` as the GPU arm prints for this workload (its step is the full slab; this
    # arm's step is the bounded sample named in cpu_baseline.sample)
    st, nd, _, nf, npx, _ = synthetic.SHAPES[args.workload]
    if args.workload == "skalow":
        nf = 1
    slots = args.slots or auto_slots(args.workload)
    return {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_s / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference",
            "config": workload_config(args.workload, npx, nd, slots * nf * st, slots),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample + ", per step"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def workload_config(workload, n, n_d, cols, slots):
    names = {"lofar": "synthetic LOFAR-HBA shape: 62 stations, 45 directions, 24 freqs, 512x512 grid, KL + Voronoi",
             "voronoi1k": "synthetic Voronoi with Gaussian smoothing at 1024x1024 grid, 62 stations, 45 directions, 24 freqs",
             "skalow": "synthetic SKA-Low shape: 512 stations, 100 directions, 2048x2048 grid, KL (one freq per step)",
             "small": "small synthetic test shape"}
    cfg = {"workload": names[workload], "image": f"{n}x{n}", "directions": n_d}
    if cols is not None:
        cfg.update({"time_slots_per_step": slots, "columns_per_step": cols,
                    "cache": f"each step writes {cols * n * n * 16 / 1e9:.1f} GB of cube into a 2-slab "
                             "ring (far larger than the 126 MB L2), no explicit flush needed",
                    "sharding": "time slabs per rank, no data-path collective"})
    return cfg


def block(res, key, world, steps):
    """One result block of a measured workload as it appears in the JSON line."""
    o = res["out"][key]
    ms = o["ms"]
    b = {"value": world * steps * res["voxels_per_step"] / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps,
         "roofline": o["roofline"], "gpu_launches": o["launches"], "clocks": o["clocks"]}
    if "smooth_pix" in o:
        b["smooth_pix"] = o["smooth_pix"]
    if "fit" in o:
        b["fit"] = o["fit"]
    return b


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
        world = res["world"]
        hb = block(res, head, world, args.steps)
        line = {"metric": METRIC, "value": hb["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": hb["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "path": "KL (fit + evaluation)" if head == "kl" else "Voronoi (fill + smoothing)",
                "config": workload_config(args.workload, res["n"], res["n_d"], res["cols"], res["slots"]),
                "roofline": hb["roofline"], "gpu_launches": hb["gpu_launches"], "clocks": hb["clocks"],
                "setup_s": out["setup"]}
        if head == "kl":
            line["fit"] = out["kl"]["fit"]
            line["fit_host_ms_per_slab"] = out["kl"]["fit_host_ms"]
            line["fit_slabs_in_timed_region"] = out["kl"]["fit_slabs"]
            line["fit_breakdown"] = out["kl"]["fit_breakdown"]
        names = {"voronoi": "voronoi", "gain": "kl_gain_eval", "kl_pairs": "kl_plane_pairs_big_endian",
                 "voronoi_direct": "voronoi_direct_smoothing"}
        for key, label in names.items():
            if key in out and key != head:
                line[label] = block(res, key, world, args.steps)
        if res["extra"]:
            line["extra"] = {}
            for name, sub in res["extra"].items():
                sub_out = sub["out"]
                sub_head = "kl" if "kl" in sub_out else "voronoi"
                eb = block(sub, sub_head, world, sub["steps"])
                eb.update(config=workload_config(name, sub["n"], sub["n_d"], sub["cols"], sub["slots"]),
                          steps=sub["steps"], warmup=sub["warmup"], setup_s=sub_out["setup"],
                          path="KL (fit + evaluation)" if sub_head == "kl" else "Voronoi (smoothed, linear form)")
                for key, label in names.items():
                    if key in sub_out and key != sub_head:
                        eb[label] = block(sub, key, world, sub["steps"])
                line["extra"][name] = eb
        if res["e2e"] is not None:
            line["e2e"] = res["e2e"]
        if world == 1 and not args.no_cpu_baseline:
            restore_affinity()
            ncpu = os.cpu_count() or 1
            value, secs, kind, cores, sample = cpu_arm(args.workload, res["smooth_pix"], ncpu)
            line["cpu_baseline"] = {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "seconds": secs,
                                    "sample": sample}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

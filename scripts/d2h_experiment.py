"""Where does the end-to-end figure go?  Separates the PCIe link from the host side.

Run on 1..N GPUs of one box (one process per GPU under torchrun, or alone):

    python scripts/d2h_experiment.py [out.txt]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29541 scripts/d2h_experiment.py out.txt

Per rank, all ranks at once (barrier before every measurement), device-timed with CUDA events:
  A. D2H copy bandwidth into pinned host memory for 64 MB .. 4 GB copies,
     cudaHostAlloc(default) vs cudaHostAlloc(portable) vs cudaHostRegister(malloc),
     with the rank's threads bound to the GPU's NUMA-local CPUs and unbound.
Rank 0 alone, host only:
  B. pwrite into a file on /dev/shm (tmpfs) with 1..32 threads, fresh pages and rewritten pages;
  C. memcpy (numpy) between host buffers with 1..16 threads.
Prints one table; bench.py's e2e numbers are explained by the smaller of A and B.
"""
import ctypes
import os
import sys
import threading
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def numa_cpus(index):
    try:
        import pynvml

        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(visible.split(",")[index]) if visible and visible.replace(",", "").isdigit() else index
        handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        return {64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1}
    except Exception:  # noqa: BLE001
        return set()


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
    rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
    rt.cudaHostRegister.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint]
    rt.cudaHostUnregister.argtypes = [ctypes.c_void_p]
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    lines = []

    def say(text):
        if rank == 0:
            print(text, flush=True)
            lines.append(text)

    all_cpus = os.sched_getaffinity(0)
    local_cpus = numa_cpus(local) & all_cpus
    say(f"# box: {torch.cuda.get_device_name(local)}, {world} rank(s), {os.cpu_count()} host CPUs visible "
        f"({len(all_cpus)} in the affinity mask), GPU {local} NUMA-local CPUs: {len(local_cpus)}")
    try:
        with open("/proc/meminfo", encoding="utf8") as fh:
            mem = {ln.split(":")[0]: ln.split(":")[1].strip() for ln in fh if ":" in ln}
        say(f"# MemTotal {mem.get('MemTotal')}, MemAvailable {mem.get('MemAvailable')}, Hugepagesize {mem.get('Hugepagesize')}")
        st = os.statvfs("/dev/shm")
        say(f"# /dev/shm: {st.f_bavail * st.f_frsize / 1e9:.0f} GB free of {st.f_blocks * st.f_frsize / 1e9:.0f} GB")
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")]
        say(f"# NUMA nodes: {len(nodes)}")
    except OSError as exc:
        say(f"# (no /proc/meminfo or /dev/shm: {exc})")

    big = 4 << 30
    src = torch.empty(big, dtype=torch.uint8, device="cuda")
    src.random_(0, 255)
    stream = torch.cuda.current_stream().cuda_stream

    def d2h_rate(host_ptr, nbytes, reps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        rt.cudaMemcpyAsync(host_ptr, src.data_ptr(), nbytes, 2, stream)       # warm
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(reps):
            rt.cudaMemcpyAsync(host_ptr, src.data_ptr(), nbytes, 2, stream)
        ev1.record()
        torch.cuda.synchronize()
        gbps = reps * nbytes / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
        if dist is not None:
            t = torch.tensor([gbps], dtype=torch.float64, device="cuda")
            lo = t.clone()
            dist.all_reduce(t)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            return float(t.item()), float(lo.item())
        return gbps, gbps

    say("\n## A. D2H into pinned host memory, all ranks at once: aggregate GB/s (slowest rank GB/s)")
    say(f"{'binding':<10}{'allocation':<28}" + "".join(f"{s >> 20:>10} MB" for s in (64 << 20, 512 << 20, 2 << 30, 4 << 30)))
    for binding in ("numa", "none"):
        if binding == "numa" and local_cpus:
            os.sched_setaffinity(0, local_cpus)
        else:
            os.sched_setaffinity(0, all_cpus)
        for alloc in ("cudaHostAlloc default", "cudaHostAlloc portable", "cudaHostRegister(malloc)"):
            ptr = ctypes.c_void_p()
            keep = None
            t0 = time.perf_counter()
            if alloc.startswith("cudaHostAlloc"):
                rc = rt.cudaHostAlloc(ctypes.byref(ptr), big, 1 if "portable" in alloc else 0)
            else:
                keep = np.empty(big + 4096, np.uint8)
                keep[::4096] = 0                                              # first touch under this binding
                addr = (keep.ctypes.data + 4095) & ~4095
                ptr = ctypes.c_void_p(addr)
                rc = rt.cudaHostRegister(ptr, big, 0)
            t_alloc = time.perf_counter() - t0
            if rc != 0:
                say(f"{binding:<10}{alloc:<28} failed with cuda error {rc}")
                continue
            cells = []
            for nbytes in (64 << 20, 512 << 20, 2 << 30, 4 << 30):
                tot, lo = d2h_rate(ptr, nbytes, max(2, (8 << 30) // nbytes))
                cells.append(f"{tot:7.1f} ({lo:5.1f})")
            say(f"{binding:<10}{alloc:<28}" + "".join(f"{c:>13}" for c in cells) + f"   [pinning 4 GiB took {t_alloc:.2f} s]")
            if alloc.startswith("cudaHostAlloc"):
                rt.cudaFreeHost(ptr)
            else:
                rt.cudaHostUnregister(ptr)
            del keep
    os.sched_setaffinity(0, all_cpus)
    del src
    if dist is not None:
        dist.barrier()

    if rank == 0:
        say("\n## B. pwrite into /dev/shm (rank 0 alone): GB/s, 8 GiB file, 64 MiB per call")
        size, chunk = 8 << 30, 64 << 20
        buf = np.random.default_rng(0).integers(0, 255, chunk, dtype=np.uint8)
        path = "/dev/shm/ssf_d2h_experiment.bin"
        try:
            say(f"{'threads':<10}{'fresh pages':>14}{'rewritten pages':>18}")
            for threads in (1, 2, 4, 8, 16, 32):
                if threads > 2 * (os.cpu_count() or 1):
                    break
                rates = []
                for fresh in (True, False):
                    if fresh and os.path.exists(path):
                        os.unlink(path)
                    fd = os.open(path, os.O_WRONLY | os.O_CREAT, 0o600)
                    os.ftruncate(fd, size)
                    offs = list(range(0, size, chunk))
                    t0 = time.perf_counter()
                    with ThreadPoolExecutor(threads) as pool:
                        list(pool.map(lambda o: os.pwrite(fd, buf, o), offs))
                    rates.append(size / (time.perf_counter() - t0) / 1e9)
                    os.close(fd)
                say(f"{threads:<10}{rates[0]:>14.1f}{rates[1]:>18.1f}")
        except OSError as exc:
            say(f"pwrite test failed: {exc}")
        finally:
            if os.path.exists(path):
                os.unlink(path)
        say("\n## C. host memcpy (numpy copyto, rank 0 alone): GB/s copied (read + write = 2x that), 4 GiB")
        a = np.ones(4 << 30, np.uint8)
        b = np.empty_like(a)
        b[::4096] = 0
        say(f"{'threads':<10}{'GB/s':>10}")
        for threads in (1, 2, 4, 8, 16):
            if threads > (os.cpu_count() or 1):
                break
            parts = [(i * len(a) // threads, (i + 1) * len(a) // threads) for i in range(threads)]
            t0 = time.perf_counter()
            ths = [threading.Thread(target=lambda lo, hi: np.copyto(b[lo:hi], a[lo:hi]), args=p) for p in parts]
            for t in ths:
                t.start()
            for t in ths:
                t.join()
            say(f"{threads:<10}{len(a) / (time.perf_counter() - t0) / 1e9:>10.1f}")
        if out_path:
            os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
            with open(out_path, "w", encoding="utf8") as fh:
                fh.write("\n".join(lines) + "\n")
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

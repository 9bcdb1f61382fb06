"""How fast can finished cube slabs be placed into a FITS file on this host?  Host only.

Screen.write's last stage copies a staged slab (pinned host memory) into the output file at its
byte offset, each plane pair twice for phase-only cubes.  Round 2's first run measured 3.5 GB/s
for 16 threads of pwritev into ONE tmpfs file -- writes to one inode serialise on its lock.  This
script times the alternatives on the box, fresh pages and already-allocated pages:

  pwrite    T threads, one fd, disjoint ranges              (what write_columns did)
  mmap      T threads copying into disjoint ranges of one shared mapping (page faults run in parallel)
  mmap+fallocate   pages allocated first by one fallocate call, then the same copies
  native    the library's ssf_host_place (std::thread, streaming stores) if it is built

    python scripts/file_sink_experiment.py [out.txt] [GiB]
"""
import ctypes
import mmap
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else None
    gib = float(sys.argv[2]) if len(sys.argv) > 2 else 4.0
    nbytes = int(gib * (1 << 30)) // (1 << 22) * (1 << 22)
    lines = []

    def say(s):
        print(s, flush=True)
        lines.append(s)

    say(f"# file sink experiment: {nbytes / 2**30:.2f} GiB per trial, cpus {os.cpu_count()}, "
        f"affinity {len(os.sched_getaffinity(0))}")
    for p in ("/sys/kernel/mm/transparent_hugepage/shmem_enabled", "/sys/kernel/mm/transparent_hugepage/enabled"):
        try:
            say(f"# {p}: {open(p).read().strip()}")
        except OSError as exc:
            say(f"# {p}: {exc}")
    try:
        st = os.statvfs("/dev/shm")
        say(f"# /dev/shm free {st.f_bavail * st.f_frsize / 2**30:.1f} GiB")
    except OSError:
        pass
    src = np.empty(nbytes, np.uint8)
    src[:] = 7
    path = "/dev/shm/ssf_sink_experiment.bin"

    def fresh():
        if os.path.exists(path):
            os.unlink(path)
        fd = os.open(path, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
        os.ftruncate(fd, nbytes)
        return fd

    def bounds(t):
        edges = [nbytes * i // t // 4096 * 4096 for i in range(t)] + [nbytes]
        return list(zip(edges[:-1], edges[1:]))

    view = memoryview(src)

    def run_pwrite(fd, t):
        def work(lo, hi):
            pos = lo
            while pos < hi:
                pos += os.pwrite(fd, view[pos:min(hi, pos + (1 << 28))], pos)
        with ThreadPoolExecutor(t) as pool:
            t0 = time.perf_counter()
            for f in [pool.submit(work, lo, hi) for lo, hi in bounds(t)]:
                f.result()
            return time.perf_counter() - t0

    def run_mmap(fd, t):
        mm = mmap.mmap(fd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        dst = np.frombuffer(mm, np.uint8)

        def work(lo, hi):
            step = 1 << 26
            for p in range(lo, hi, step):
                q = min(hi, p + step)
                np.copyto(dst[p:q], src[p:q])
        with ThreadPoolExecutor(t) as pool:
            t0 = time.perf_counter()
            for f in [pool.submit(work, lo, hi) for lo, hi in bounds(t)]:
                f.result()
            dt = time.perf_counter() - t0
        del dst
        mm.close()
        return dt

    lib = None
    try:
        from ska_sdp_screen_fitting_b200 import _lib
        lib = _lib.load()
        if not hasattr(lib, "ssf_host_place"):
            lib = None
    except Exception as exc:  # noqa: BLE001
        say(f"# native library not available: {exc}")

    def run_native(fd, t):
        mm = mmap.mmap(fd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        dst = np.frombuffer(mm, np.uint8)
        t0 = time.perf_counter()
        # one "column" = 4 MiB source placed once
        rc = lib.ssf_host_place(ctypes.c_void_p(dst.ctypes.data), ctypes.c_void_p(src.ctypes.data),
                                ctypes.c_int64(nbytes >> 22), ctypes.c_int64(1 << 22), ctypes.c_int(1),
                                ctypes.c_int(t))
        dt = time.perf_counter() - t0
        assert rc == 0
        del dst
        mm.close()
        return dt

    say("method              threads  fresh GB/s  rewrite GB/s")
    threads = [1, 2, 4, 8, 16, 32]
    for name, fn in (("pwrite", run_pwrite), ("mmap copy", run_mmap), ("native place", run_native)):
        if fn is run_native and lib is None:
            continue
        for t in threads:
            fd = fresh()
            a = fn(fd, t)
            b = fn(fd, t)
            os.close(fd)
            say(f"{name:18s}  {t:7d}  {nbytes / a / 1e9:10.2f}  {nbytes / b / 1e9:12.2f}")
    # fallocate first (one call allocates and zeroes all pages under the inode lock)
    fd = fresh()
    t0 = time.perf_counter()
    os.posix_fallocate(fd, 0, nbytes)
    dt = time.perf_counter() - t0
    os.close(fd)
    say(f"posix_fallocate of the whole file: {nbytes / dt / 1e9:.2f} GB/s")
    # page-fault rate alone: threads touching one byte per page of a fresh mapping
    for t in (1, 4, 16):
        fd = fresh()
        mm = mmap.mmap(fd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        dst = np.frombuffer(mm, np.uint8)

        def touch(lo, hi):
            dst[lo:hi:4096] = 1
        with ThreadPoolExecutor(t) as pool:
            t0 = time.perf_counter()
            for f in [pool.submit(touch, lo, hi) for lo, hi in bounds(t)]:
                f.result()
            dt = time.perf_counter() - t0
        del dst
        mm.close()
        os.close(fd)
        say(f"first touch only (1 B per 4 KiB page), {t:2d} threads: {nbytes / dt / 1e9:.2f} GB/s of pages")
    # anonymous memory for comparison (what a pinned staging buffer costs nothing of)
    for t in (1, 16):
        dst = np.empty(nbytes, np.uint8)

        def work(lo, hi):
            np.copyto(dst[lo:hi], src[lo:hi])
        with ThreadPoolExecutor(t) as pool:
            t0 = time.perf_counter()
            for f in [pool.submit(work, lo, hi) for lo, hi in bounds(t)]:
                f.result()
            a = time.perf_counter() - t0
            t0 = time.perf_counter()
            for f in [pool.submit(work, lo, hi) for lo, hi in bounds(t)]:
                f.result()
            b = time.perf_counter() - t0
        say(f"anonymous memcpy, {t:2d} threads: fresh {nbytes / a / 1e9:.2f} GB/s, touched {nbytes / b / 1e9:.2f} GB/s")
        del dst
    if os.path.exists(path):
        os.unlink(path)
    if out_path:
        os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
        with open(out_path, "w", encoding="utf8") as fh:
            fh.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()

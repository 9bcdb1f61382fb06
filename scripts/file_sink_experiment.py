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
    # ---- combinations: who allocates the pages, who maps them, who copies
    libc = ctypes.CDLL(None, use_errno=True)
    libc.fallocate.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int64, ctypes.c_int64]
    libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    MADV_POPULATE_WRITE = 23

    def mapped(fd):
        mm = mmap.mmap(fd, nbytes, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
        return mm, np.frombuffer(mm, np.uint8)

    def copy_threads(dst, t, gate=None):
        def work(lo, hi):
            step = 1 << 25
            for p in range(lo, hi, step):
                q = min(hi, p + step)
                if gate is not None:
                    gate(q)
                np.copyto(dst[p:q], src[p:q])
        with ThreadPoolExecutor(t) as pool:
            for f in [pool.submit(work, lo, hi) for lo, hi in bounds(t)]:
                f.result()

    import threading
    for t in (8, 16):
        # (a) one fallocate of everything, then copies
        fd = fresh()
        t0 = time.perf_counter()
        libc.fallocate(fd, 0, 0, nbytes)
        t1 = time.perf_counter()
        mm, dst = mapped(fd)
        copy_threads(dst, t)
        t2 = time.perf_counter()
        del dst
        mm.close()
        os.close(fd)
        say(f"fallocate then copy, {t:2d} threads: fallocate {nbytes / (t1 - t0) / 1e9:.2f} GB/s, copy "
            f"{nbytes / (t2 - t1) / 1e9:.2f} GB/s, together {nbytes / (t2 - t0) / 1e9:.2f} GB/s")
        # (b) fallocate thread running ahead in 64 MiB calls, copies in SEQUENTIAL 256 MiB slabs behind it
        fd = fresh()
        front = [0]
        cv = threading.Condition()

        def allocator():
            pos = 0
            while pos < nbytes:
                n = min(64 << 20, nbytes - pos)
                libc.fallocate(fd, 0, pos, n)
                pos += n
                with cv:
                    front[0] = pos
                    cv.notify_all()
        mm, dst = mapped(fd)
        t0 = time.perf_counter()
        th = threading.Thread(target=allocator)
        th.start()
        slab = 256 << 20
        with ThreadPoolExecutor(t) as pool:
            for s0 in range(0, nbytes, slab):
                s1 = min(nbytes, s0 + slab)
                with cv:
                    while front[0] < s1:
                        cv.wait()
                edges = [s0 + (s1 - s0) * i // t // 4096 * 4096 for i in range(t)] + [s1]
                for f in [pool.submit(np.copyto, dst[a:b], src[a:b]) for a, b in zip(edges[:-1], edges[1:])]:
                    f.result()
        th.join()
        t1 = time.perf_counter()
        del dst
        mm.close()
        os.close(fd)
        say(f"fallocate thread ahead + slab copies, {t:2d} threads: {nbytes / (t1 - t0) / 1e9:.2f} GB/s")
        # (c) MADV_POPULATE_WRITE by the copy threads on their own 32 MiB pieces, then the copy
        fd = fresh()
        mm, dst = mapped(fd)
        base = dst.ctypes.data

        def populate(q_unused, _state={}):
            return None
        def work_pop(lo, hi):
            step = 1 << 25
            for p in range(lo, hi, step):
                q = min(hi, p + step)
                libc.madvise(ctypes.c_void_p(base + p), q - p, MADV_POPULATE_WRITE)
                np.copyto(dst[p:q], src[p:q])
        t0 = time.perf_counter()
        with ThreadPoolExecutor(t) as pool:
            for f in [pool.submit(work_pop, lo, hi) for lo, hi in bounds(t)]:
                f.result()
        t1 = time.perf_counter()
        del dst
        mm.close()
        os.close(fd)
        say(f"madvise(POPULATE_WRITE) + copy per 32 MiB piece, {t:2d} threads: {nbytes / (t1 - t0) / 1e9:.2f} GB/s")
        # (d) fallocate of everything, then madvise populate + copy
        fd = fresh()
        t0 = time.perf_counter()
        libc.fallocate(fd, 0, 0, nbytes)
        mm, dst = mapped(fd)
        base = dst.ctypes.data
        with ThreadPoolExecutor(t) as pool:
            for f in [pool.submit(work_pop, lo, hi) for lo, hi in bounds(t)]:
                f.result()
        t1 = time.perf_counter()
        del dst
        mm.close()
        os.close(fd)
        say(f"fallocate, then madvise(POPULATE_WRITE) + copy, {t:2d} threads: together {nbytes / (t1 - t0) / 1e9:.2f} GB/s")
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

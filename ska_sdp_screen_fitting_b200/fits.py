"""
Minimal FITS primary-HDU writer / reader for the a-term cube (no astropy / cfitsio here).

Layout written: one primary HDU, BITPIX -32, big-endian float32, numpy shape
``[time, freq, ant, 4, y, x]`` (processing_utils.py:190-194, 202), header cards in the order
``make_template_image`` creates them (processing_utils.py:204-289).

The writer streams: ``CubeWriter`` creates the file with its full header, and time slabs are
written at their byte offset as they come off the GPU, so a cube never has to exist in host
memory and every byte is written once (the reference writes the cube twice,
screen.py:331-334 then :381-382).
"""
import ctypes
import mmap
import os
import threading
import time

import numpy as np

BLOCK = 2880
try:
    _IOV_MAX = min(1024, os.sysconf("SC_IOV_MAX"))
except (ValueError, OSError, AttributeError):        # pragma: no cover
    _IOV_MAX = 1024


def _fmt_float(value):
    """astropy.io.fits.card._format_float: 16 significant digits, always a decimal point."""
    text = f"{float(value):.16G}"
    if "." not in text and "E" not in text:
        text += ".0"
    elif "E" in text:
        mant, exp = text.split("E")
        sign = "-" if exp[0] == "-" else ""
        exp = exp.lstrip("+-").lstrip("0")
        if "." not in mant:
            mant += ".0"
        text = f"{mant}E{sign}{int(exp or 0):02d}"
    return text[:20]


def _card(key, value, comment=""):
    if isinstance(value, (bool, np.bool_)):
        val = f"{'T' if value else 'F':>20}"
    elif isinstance(value, (int, np.integer)):
        val = f"{int(value):>20d}"
    elif isinstance(value, (float, np.floating)):
        val = f"{_fmt_float(value):>20}"
    elif isinstance(value, str):
        val = "'" + f"{value.replace(chr(39), chr(39) * 2):<8}" + "'"
        val = f"{val:<20}"
    else:
        raise TypeError(f"unsupported FITS value {value!r}")
    text = f"{key:<8}= {val}"
    if comment:
        text += f" / {comment}"
    if len(text) > 80:
        text = text[:80]
    return f"{text:<80}"


def header_bytes(cards):
    text = "".join(_card(*c) for c in cards) + f"{'END':<80}"
    pad = (-len(text)) % BLOCK
    return (text + " " * pad).encode("ascii")


def aterm_cards(shape, rad, dec, cellsize_deg, freqs, times, aterm_type="gain"):
    """Header of an a-term cube (processing_utils.py:202-289).  ``shape`` is the numpy shape
    [nt, nf, na, 4, ny, nx] for ``aterm_type="gain"`` (what Screen.write produces,
    screen.py:182-190) or [nt, nf, na, ny, nx] for ``"tec"`` (no MATRIX axis, :187-190,
    :230-238)."""
    if aterm_type not in ("gain", "tec"):
        raise ValueError("aterm_type must be 'gain' or 'tec'")
    if len(shape) != (6 if aterm_type == "gain" else 5):
        raise ValueError(f"a {aterm_type} cube has {6 if aterm_type == 'gain' else 5} axes, got shape {tuple(shape)}")
    nt, nf = shape[0], shape[1]
    ny, nx = shape[-2], shape[-1]
    freqs = np.asarray(freqs, np.float64)
    times = np.asarray(times, np.float64)
    cards = [("SIMPLE", True, "conforms to FITS standard"),
             ("BITPIX", -32, "array data type"),
             ("NAXIS", len(shape), "number of array dimensions")]
    cards += [(f"NAXIS{i + 1}", int(n)) for i, n in enumerate(reversed(shape))]
    cards.append(("EXTEND", True))

    def axis(i, crval, cdelt, crpix, cunit, ctype):
        return [(f"CRVAL{i}", crval), (f"CDELT{i}", cdelt), (f"CRPIX{i}", crpix),
                (f"CUNIT{i}", cunit), (f"CTYPE{i}", ctype)]

    cards += axis(1, float(rad), -float(cellsize_deg), nx / 2.0, "deg", "RA---SIN")
    cards += axis(2, float(dec), float(cellsize_deg), ny / 2.0, "deg", "DEC--SIN")
    i = 3
    if aterm_type == "gain":
        cards += axis(i, 0.0, 1.0, 1.0, "", "MATRIX")
        i += 1
    cards += axis(i, 0.0, 1.0, 1.0, "", "ANTENNA")
    del_freq = float(np.min(freqs[1:] - freqs[:-1])) if nf > 1 else 1e8       # :249-253
    cards.append(("RESTFRQ", float(freqs[0])))
    cards += axis(i + 1, float(freqs[0]), del_freq, 1.0, "Hz", "FREQ")
    if nt > 1:                                                                 # :263-276
        deltas = times[1:] - times[:-1]
        del_time = float(np.min(deltas[:-1])) if nt > 2 else float(deltas[0])
    else:
        del_time = 1.0
    cards += axis(i + 2, float(times[0]), del_time, 1.0, "s", "TIME")
    cards += [("EQUINOX", 2000.0), ("TELESCOP", "LOFAR")]
    return cards


class CubeWriter:
    """Creates ``path`` with its header and accepts time slabs in any order."""

    def __init__(self, path, shape, cards, create=True):
        """``create=False`` opens a file another process (rank 0) has already laid out."""
        self.path = path
        self.shape = tuple(int(s) for s in shape)
        self._head = header_bytes(cards)
        self._row_bytes = int(np.prod(self.shape[1:])) * 4
        data_bytes = self.shape[0] * self._row_bytes
        self._total = len(self._head) + data_bytes + (-data_bytes) % BLOCK
        self._map = None
        self._map_view = None
        self._alloc_thread = None
        self._alloc_front = 0                        # bytes [_alloc_lo, _alloc_front) have their pages
        self._alloc_lo = 0
        self._alloc_stop = False
        self._alloc_cv = threading.Condition()
        if create:
            self._fd = os.open(path, os.O_RDWR | os.O_CREAT | os.O_TRUNC, 0o644)
            os.pwrite(self._fd, self._head, 0)
            os.ftruncate(self._fd, self._total)      # zero padding to a whole block
        else:
            self._fd = os.open(path, os.O_RDWR)

    def mapped_address(self):
        """Address of a shared read-write mapping of the whole file (made on first use), or None
        where the file system cannot map it.  Copies into the mapping from several threads fault
        the pages of a fresh tmpfs / page-cache file in parallel; ``pwrite`` calls on one file
        take turns on the inode lock (measured: profiles/r02_file_sink.txt)."""
        if self._map is None:
            try:
                self._map = mmap.mmap(self._fd, self._total, mmap.MAP_SHARED, mmap.PROT_READ | mmap.PROT_WRITE)
            except (OSError, ValueError):
                self._map = False
                return None
            self._map_view = np.frombuffer(self._map, np.uint8)
        return int(self._map_view.ctypes.data) if self._map else None

    def preallocate(self, c0, ncol, chunk=64 << 20):
        """Starts a background thread that allocates the pages of columns [c0, c0 + ncol) ahead of
        the writers, ``fallocate`` call by call.  On a tmpfs / page-cache file the first touch of
        a page is the expensive part of placing a slab: measured on the GPU boxes
        (profiles/r02_file_sink.txt), threads faulting pages in through a mapping top out at
        ~8 GB/s however many they are, one fallocate stream allocates 15 GB/s, and copies into
        pages that exist run at 30-40 GB/s -- but the two together (one after the other, or the
        allocator running ahead) still end at 8-10 GB/s, so Screen.write leaves this off by
        default.  Silently does nothing where the file system has no fallocate (the raw system
        call: glibc's posix_fallocate emulation would race with the writers)."""
        col_bytes = int(np.prod(self.shape[3:])) * 4
        lo = len(self._head) + c0 * col_bytes
        hi = lo + ncol * col_bytes
        try:
            libc = ctypes.CDLL(None, use_errno=True)
            fallocate = libc.fallocate
            fallocate.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int64, ctypes.c_int64]
            fallocate.restype = ctypes.c_int
        except (OSError, AttributeError):
            return
        self._alloc_lo = self._alloc_front = lo
        fd = self._fd

        def run():
            pos = lo
            while pos < hi and not self._alloc_stop:
                n = min(chunk, hi - pos)
                if fallocate(fd, 0, pos, n) != 0:       # not supported / no space: the writers will fault
                    pos = hi
                else:
                    pos += n
                with self._alloc_cv:
                    self._alloc_front = pos
                    self._alloc_cv.notify_all()
            with self._alloc_cv:
                self._alloc_front = hi
                self._alloc_cv.notify_all()

        self._alloc_thread = threading.Thread(target=run, name="ssf-fallocate", daemon=True)
        self._alloc_thread.start()

    def _wait_allocated(self, end):
        if self._alloc_thread is None:
            return
        with self._alloc_cv:
            while self._alloc_front < end and self._alloc_thread.is_alive():
                self._alloc_cv.wait(0.05)

    def write_times(self, t0, slab, already_big_endian=False):
        """slab: array whose leading axis is time, starting at time index ``t0``."""
        arr = np.ascontiguousarray(slab)
        if arr.dtype.itemsize != 4:
            raise TypeError("slab must be a 4-byte type")
        if not already_big_endian:
            arr = arr.astype(">f4", copy=False)
        nbytes = arr.nbytes
        if nbytes % self._row_bytes or t0 * self._row_bytes + nbytes > self.shape[0] * self._row_bytes:
            raise ValueError("slab does not fit the cube")
        view = memoryview(arr).cast("B")
        off = len(self._head) + t0 * self._row_bytes
        done = 0
        while done < nbytes:                         # pwrite caps at ~2 GiB per call
            done += os.pwrite(self._fd, view[done:done + (1 << 30)], off + done)

    def write_columns(self, c0, slab, already_big_endian=False, pairs=False, pool=None, place=None):
        """slab: whole (time, freq, antenna) columns -- [ncol, 4, ny, nx] -- starting at the flat
        column index ``c0`` = (t * n_freq + f) * n_ant + a.  The cube is column-major in exactly
        that order (processing_utils.py:190-194), so any run of columns is one byte range.

        ``pairs=True``: slab holds only the first half of every column ([ncol, 2, ny, nx]: planes
        0, 1 of a phase-only cube, whose planes 2, 3 repeat them -- kl_screen.py:368-373); each
        half is placed twice with one gathered write (``pwritev``), so the copy that crossed
        PCIe is half the file.  ``pool``: a ThreadPoolExecutor; the columns are split among its
        workers (``pwrite`` releases the GIL; one thread fills a tmpfs / page cache at only a few
        GB/s).  ``place``: ``callable(dst_address, src_address, ncol, src_bytes, copies)`` -- the
        library's multi-threaded ``ssf_host_place``; used with a shared mapping of the file
        instead of the write calls wherever the file can be mapped.  Returns the seconds spent
        writing."""
        arr = np.ascontiguousarray(slab)
        if arr.dtype.itemsize != 4:
            raise TypeError("slab must be a 4-byte type")
        if not already_big_endian:
            arr = arr.astype(">f4", copy=False)
        col_bytes = int(np.prod(self.shape[3:])) * 4
        src_bytes = col_bytes // 2 if pairs else col_bytes
        nbytes = arr.nbytes
        if pairs and col_bytes % 2:
            raise ValueError("pairs need an even number of planes")
        if nbytes % src_bytes or (c0 + nbytes // src_bytes) * col_bytes > self.shape[0] * self._row_bytes:
            raise ValueError("slab does not fit the cube")
        ncol = nbytes // src_bytes
        view = memoryview(arr).cast("B")
        base = len(self._head)
        fd = self._fd
        if place is not None:
            address = self.mapped_address()
            if address is not None:
                t_begin = time.perf_counter()
                self._wait_allocated(base + (c0 + ncol) * col_bytes)
                place(address + base + c0 * col_bytes, int(arr.ctypes.data), ncol, src_bytes, 2 if pairs else 1)
                return time.perf_counter() - t_begin

        def run(lo, hi):
            if not pairs:
                off, pos, end = base + (c0 + lo) * col_bytes, lo * col_bytes, hi * col_bytes
                while pos < end:                         # pwrite caps at ~2 GiB per call
                    done = os.pwrite(fd, view[pos:min(end, pos + (1 << 30))], off)
                    pos += done
                    off += done
                return
            per_call = max(1, min(_IOV_MAX // 2, (1 << 30) // col_bytes))
            for b0 in range(lo, hi, per_call):
                b1 = min(hi, b0 + per_call)
                iov = []
                for c in range(b0, b1):
                    half = view[c * src_bytes:(c + 1) * src_bytes]
                    iov += [half, half]
                off, want = base + (c0 + b0) * col_bytes, (b1 - b0) * col_bytes
                done = os.pwritev(fd, iov, off)
                while done < want:                       # short write: finish byte by byte range
                    c, within = divmod(done, col_bytes)
                    src = (b0 + c) * src_bytes + within % src_bytes
                    left = src_bytes - within % src_bytes
                    done += os.pwrite(fd, view[src:src + left], off + done)

        t_begin = time.perf_counter()
        workers = getattr(pool, "_max_workers", 1) if pool is not None else 1
        parts = max(1, min(workers, ncol))
        if parts == 1:
            run(0, ncol)
        else:
            bounds = [ncol * i // parts for i in range(parts + 1)]
            for fut in [pool.submit(run, bounds[i], bounds[i + 1]) for i in range(parts)]:
                fut.result()
        return time.perf_counter() - t_begin

    def close(self):
        if self._alloc_thread is not None:
            self._alloc_stop = True
            self._alloc_thread.join()
            self._alloc_thread = None
        if self._map:
            self._map_view = None
            self._map.close()
        self._map = None
        if self._fd is not None:
            os.close(self._fd)
            self._fd = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def write_image(path, data, cards):
    """Small convenience: whole array in one go (used for the Voronoi template image)."""
    data = np.asarray(data, np.float32)
    with CubeWriter(path, data.shape, cards) as wr:
        wr.write_times(0, data)


def read(path, memmap=True):
    """Returns (header dict in card order, data array native-endian view or memmap)."""
    header = {}
    with open(path, "rb") as fh:
        pos = 0
        done = False
        while not done:
            block = fh.read(BLOCK)
            if len(block) < BLOCK:
                raise ValueError("truncated FITS header")
            pos += BLOCK
            for i in range(0, BLOCK, 80):
                card = block[i:i + 80].decode("ascii")
                key = card[:8].strip()
                if key == "END":
                    done = True
                    break
                if card[8:10] != "= ":
                    continue
                raw = card[10:].split(" / ")[0].strip() if not card[10:].lstrip().startswith("'") \
                    else card[10:]
                if raw.lstrip().startswith("'"):
                    body = raw.lstrip()[1:]
                    end = body.find("'")
                    while end + 1 < len(body) and body[end + 1] == "'":
                        end = body.find("'", end + 2)
                    header[key] = body[:end].replace("''", "'").rstrip()
                elif raw in ("T", "F"):
                    header[key] = raw == "T"
                else:
                    try:
                        header[key] = int(raw)
                    except ValueError:
                        header[key] = float(raw)
    naxis = header["NAXIS"]
    shape = tuple(header[f"NAXIS{i}"] for i in range(naxis, 0, -1))
    dtype = {-32: ">f4", -64: ">f8", 32: ">i4", 16: ">i2"}[header["BITPIX"]]
    if memmap:
        data = np.memmap(path, dtype=dtype, mode="r", offset=pos, shape=shape)
    else:
        data = np.fromfile(path, dtype=dtype, offset=pos, count=int(np.prod(shape))).reshape(shape)
    return header, data

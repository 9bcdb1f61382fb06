"""
Minimal read-only HDF5 reader -- just enough for DP3 / LoSoTo h5parm files.

There is no HDF5 library on this image (no h5py / pytables / libhdf5), so the read
side of the reference's ``utils/h5parm.py`` (``h5parm.py:107-120`` open,
``:731-767`` antenna / source tables, ``:1306-1322`` axis datasets) is served by this
parser of the on-disk format:

* superblock version 0/1, 8-byte (or 4-byte) offsets and lengths
* "old style" groups: symbol-table message -> v1 B-tree (``TREE``) -> ``SNOD`` -> local heap
* version-1 object headers with continuation blocks
* datasets: contiguous, compact, or chunked (v1 chunk B-tree) with optional
  deflate + shuffle filters
* datatypes: fixed-point, IEEE float, fixed-length string, compound (v1-v3), array
* attributes (message 0x0C, versions 1-3) with fixed-length or variable-length-free types

Anything else raises ``H5FormatError`` -- loudly, never a silent wrong read.
"""
import mmap
import struct
import zlib

import numpy as np

_SIG = b"\x89HDF\r\n\x1a\n"
_UNDEF = 0xFFFFFFFFFFFFFFFF


class H5FormatError(Exception):
    pass


class _Node:
    def __init__(self, f, name, addr):
        self._f = f
        self.name = name
        self.addr = addr
        self._msgs = f._read_object_header(addr)
        self._attrs = None

    @property
    def attrs(self):
        if self._attrs is None:
            self._attrs = {}
            for mtype, body in self._msgs:
                if mtype == 0x0C:
                    k, v = self._f._parse_attribute(body)
                    self._attrs[k] = v
        return self._attrs


class Group(_Node):
    def __init__(self, f, name, addr):
        super().__init__(f, name, addr)
        self._children = None

    def _links(self):
        if self._children is None:
            self._children = {}
            for mtype, body in self._msgs:
                if mtype == 0x11:
                    btree, heap = self._f._unpack_oo(body, 0)
                    self._children = self._f._read_symbol_table(btree, heap)
                    break
            else:
                if any(m == 0x02 or m == 0x06 for m, _ in self._msgs):
                    raise H5FormatError("new-style (link message) groups are not supported")
        return self._children

    def keys(self):
        return list(self._links().keys())

    def __contains__(self, key):
        return key.strip("/").split("/")[0] in self._links()

    def __getitem__(self, path):
        node = self
        for part in [p for p in path.split("/") if p]:
            links = node._links()
            if part not in links:
                raise KeyError(f"{part!r} not found in {node.name!r}")
            node = self._f._open_object(f"{node.name.rstrip('/')}/{part}", links[part])
        return node


class Dataset(_Node):
    def __init__(self, f, name, addr):
        super().__init__(f, name, addr)
        self.shape = None
        self.dtype = None
        self._layout = None
        self._filters = []
        for mtype, body in self._msgs:
            if mtype == 0x01:
                self.shape = f._parse_dataspace(body)
            elif mtype == 0x03:
                self.dtype, _ = f._parse_datatype(body, 0)
            elif mtype == 0x08:
                self._layout = f._parse_layout(body)
            elif mtype == 0x0B:
                self._filters = f._parse_filters(body)
        if self.shape is None or self.dtype is None or self._layout is None:
            raise H5FormatError(f"{name}: incomplete dataset header")

    def byte_offset(self):
        """File offset of a contiguous dataset's raw bytes (None otherwise)."""
        if self._layout[0] == "contiguous":
            return self._layout[1]
        return None

    def read(self):
        n = int(np.prod(self.shape, dtype=np.int64)) if self.shape else 1
        kind = self._layout[0]
        if kind == "contiguous":
            _, addr, _size = self._layout
            if addr == _UNDEF:
                out = np.zeros(n, dtype=self.dtype)
            else:
                out = np.frombuffer(self._f._buf, dtype=self.dtype, count=n, offset=addr).copy()
        elif kind == "compact":
            out = np.frombuffer(self._layout[1], dtype=self.dtype, count=n).copy()
        else:
            out = self._f._read_chunked(self)
        return out.reshape(self.shape)

    def read_rows(self, start, stop):
        """Rows [start, stop) of the leading axis only: a rank of a time-sharded run reads just
        its own time slots (the file is memory-mapped, so nothing else is touched; chunked
        datasets decode only the chunks that overlap the range)."""
        if not self.shape:
            raise IndexError("scalar dataset has no rows")
        start, stop = max(0, int(start)), min(int(stop), self.shape[0])
        stop = max(start, stop)
        row = int(np.prod(self.shape[1:], dtype=np.int64))
        shape = (stop - start,) + tuple(self.shape[1:])
        kind = self._layout[0]
        if kind == "contiguous":
            _, addr, _size = self._layout
            if addr == _UNDEF:
                return np.zeros(shape, dtype=self.dtype)
            return np.frombuffer(self._f._buf, dtype=self.dtype, count=(stop - start) * row,
                                 offset=addr + start * row * self.dtype.itemsize).copy().reshape(shape)
        if kind == "compact":
            return self.read()[start:stop]
        return self._f._read_chunked(self, rows=(start, stop))

    def __getitem__(self, key):
        return self.read()[key]

    def __array__(self, dtype=None, copy=None):
        a = self.read()
        return a.astype(dtype) if dtype is not None else a

    def __len__(self):
        return self.shape[0]


class File(Group):
    """``File(path)`` -> root group; ``f["sol000/phase000/val"].read()``."""

    def __init__(self, path):
        # memory-mapped, read-only: opening costs nothing and a reader touches only the pages
        # of the datasets (or rows, Dataset.read_rows) it asks for
        with open(path, "rb") as fh:
            try:
                self._buf = mmap.mmap(fh.fileno(), 0, access=mmap.ACCESS_READ)
            except (ValueError, OSError):            # empty file / no mmap on this file system
                self._buf = fh.read()
        self.filename = path
        base = 0
        while base < len(self._buf) and self._buf[base:base + 8] != _SIG:
            base = 512 if base == 0 else base * 2
        if self._buf[base:base + 8] != _SIG:
            raise H5FormatError("not an HDF5 file")
        ver = self._buf[base + 8]
        if ver not in (0, 1):
            raise H5FormatError(f"superblock version {ver} not supported (need 0 or 1)")
        self._so = self._buf[base + 13]
        self._sl = self._buf[base + 14]
        if self._so not in (4, 8) or self._sl not in (4, 8):
            raise H5FormatError("unsupported offset/length size")
        pos = base + 24 + (4 if ver == 1 else 0)
        pos += 4 * self._so  # base, free-space, eof, driver addresses
        # root symbol-table entry
        _name_off, root_addr = self._unpack_oo(self._buf, pos)
        self._cache = {}
        Group.__init__(self, self, "/", root_addr)

    def close(self):
        # arrays handed out are copies, so the mapping can go (numpy views of it would keep it alive)
        self._cache = {}
        if isinstance(self._buf, mmap.mmap):
            try:
                self._buf.close()
            except BufferError:                      # an exported view is still alive
                pass
        self._buf = b""

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False

    # ------------------------------------------------------------------ primitives
    def _uint(self, buf, pos, size):
        return int.from_bytes(buf[pos:pos + size], "little")

    def _unpack_oo(self, buf, pos):
        return self._uint(buf, pos, self._so), self._uint(buf, pos + self._so, self._so)

    def _open_object(self, name, addr):
        if addr in self._cache:
            return self._cache[addr]
        msgs = self._read_object_header(addr)
        types = {m for m, _ in msgs}
        if 0x08 in types or (0x01 in types and 0x03 in types):
            node = Dataset(self, name, addr)
        else:
            node = Group(self, name, addr)
        self._cache[addr] = node
        return node

    # ------------------------------------------------------------- object headers
    def _read_object_header(self, addr):
        buf = self._buf
        if buf[addr:addr + 4] == b"OHDR":
            raise H5FormatError("version-2 object headers are not supported")
        ver = buf[addr]
        if ver != 1:
            raise H5FormatError(f"object header version {ver} at {addr}")
        nmsg = self._uint(buf, addr + 2, 2)
        hsize = self._uint(buf, addr + 8, 4)
        blocks = [(addr + 16, hsize)]
        msgs = []
        while blocks and len(msgs) < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and len(msgs) < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", buf, pos)
                body = buf[pos + 8:pos + 8 + msize]
                pos += 8 + msize
                if mtype == 0x10:
                    off = self._uint(body, 0, self._so)
                    length = self._uint(body, self._so, self._sl)
                    blocks.append((off, length))
                msgs.append((mtype, body))
        return msgs

    # ------------------------------------------------------------------- groups
    def _heap_string(self, heap_addr, off):
        buf = self._buf
        if buf[heap_addr:heap_addr + 4] != b"HEAP":
            raise H5FormatError("bad local heap")
        data_addr = self._uint(buf, heap_addr + 8 + 2 * self._sl, self._so)
        start = data_addr + off
        end = buf.find(b"\0", start)
        return buf[start:end].decode("utf8")

    def _read_symbol_table(self, btree, heap):
        out = {}
        self._walk_group_btree(btree, heap, out)
        return out

    def _walk_group_btree(self, addr, heap, out):
        buf = self._buf
        if buf[addr:addr + 4] != b"TREE":
            raise H5FormatError("bad group B-tree node")
        level = buf[addr + 5]
        used = self._uint(buf, addr + 6, 2)
        pos = addr + 8 + 2 * self._so
        for _ in range(used):
            pos += self._sl  # key
            child = self._uint(buf, pos, self._so)
            pos += self._so
            if level > 0:
                self._walk_group_btree(child, heap, out)
            else:
                if buf[child:child + 4] != b"SNOD":
                    raise H5FormatError("bad symbol node")
                nsym = self._uint(buf, child + 6, 2)
                epos = child + 8
                for _ in range(nsym):
                    name_off, ohdr = self._unpack_oo(buf, epos)
                    out[self._heap_string(heap, name_off)] = ohdr
                    epos += 2 * self._so + 24

    # ---------------------------------------------- locations (used by utils/h5write.Appender)
    def _walk_entries(self, group):
        """Yields (name, file offset of the symbol-table entry) for every child of ``group``."""
        for mtype, body in group._msgs:
            if mtype == 0x11:
                btree, heap = self._unpack_oo(body, 0)
                break
        else:
            raise H5FormatError(f"{group.name}: not an old-style group")
        stack = [btree]
        buf = self._buf
        while stack:
            addr = stack.pop()
            level = buf[addr + 5]
            used = self._uint(buf, addr + 6, 2)
            pos = addr + 8 + 2 * self._so
            for _ in range(used):
                pos += self._sl
                child = self._uint(buf, pos, self._so)
                pos += self._so
                if level > 0:
                    stack.append(child)
                    continue
                nsym = self._uint(buf, child + 6, 2)
                for i in range(nsym):
                    epos = child + 8 + i * (2 * self._so + 24)
                    yield self._heap_string(heap, self._uint(buf, epos, self._so)), epos

    def _symbol_entries(self, group):
        """{child name: (object header address, (btree, heap) cached for groups or None)}."""
        out = {}
        for name, epos in self._walk_entries(group):
            ohdr = self._uint(self._buf, epos + self._so, self._so)
            cache = self._uint(self._buf, epos + 2 * self._so, 4)
            stab = self._unpack_oo(self._buf, epos + 2 * self._so + 8) if cache == 1 else None
            if stab is None:
                # a group whose entry carries no cached addresses: take them from its header
                for mtype, body in self._read_object_header(ohdr):
                    if mtype == 0x11:
                        stab = self._unpack_oo(body, 0)
            out[name] = (ohdr, stab)
        return out

    def _entry_offset(self, group, name):
        for child, epos in self._walk_entries(group):
            if child == name:
                return epos
        raise KeyError(name)

    def _message_offset(self, addr, want):
        """File offset of the BODY of the first message of type ``want`` in the object header."""
        buf = self._buf
        nmsg = self._uint(buf, addr + 2, 2)
        blocks = [(addr + 16, self._uint(buf, addr + 8, 4))]
        seen = 0
        while blocks and seen < nmsg:
            pos, size = blocks.pop(0)
            end = pos + size
            while pos + 8 <= end and seen < nmsg:
                mtype, msize, _flags = struct.unpack_from("<HHB", buf, pos)
                if mtype == want:
                    return pos + 8
                if mtype == 0x10:
                    blocks.append((self._uint(buf, pos + 8, self._so), self._uint(buf, pos + 8 + self._so, self._sl)))
                pos += 8 + msize
                seen += 1
        raise KeyError(f"message type {want:#x} not found")

    # ------------------------------------------------------------------ messages
    def _parse_dataspace(self, body):
        ver, rank, flags = body[0], body[1], body[2]
        if ver == 1:
            pos = 8
        elif ver == 2:
            pos = 4
            if body[3] == 2:  # null dataspace
                return (0,)
        else:
            raise H5FormatError(f"dataspace version {ver}")
        del flags
        return tuple(self._uint(body, pos + i * self._sl, self._sl) for i in range(rank))

    def _parse_datatype(self, body, pos):
        """Returns (numpy dtype, bytes consumed)."""
        cls = body[pos] & 0x0F
        ver = body[pos] >> 4
        bits0 = body[pos + 1]
        bits = body[pos + 1] | (body[pos + 2] << 8) | (body[pos + 3] << 16)
        size = self._uint(body, pos + 4, 4)
        p = pos + 8
        if cls == 0:
            order = ">" if bits0 & 1 else "<"
            signed = "i" if bits0 & 8 else "u"
            return np.dtype(f"{order}{signed}{size}"), p + 4 - pos
        if cls == 1:
            order = ">" if bits0 & 1 else "<"
            return np.dtype(f"{order}f{size}"), p + 12 - pos
        if cls == 3:
            return np.dtype(f"S{size}"), p - pos
        if cls == 6:
            nmemb = bits & 0xFFFF
            names, formats, offsets = [], [], []
            for _ in range(nmemb):
                end = body.index(b"\0", p)
                name = body[p:end].decode("utf8")
                if ver < 3:
                    p += ((end - p) // 8 + 1) * 8
                else:
                    p = end + 1
                if ver < 3:
                    off = self._uint(body, p, 4)
                    p += 4
                else:
                    nb = max(1, (max(size - 1, 1).bit_length() + 7) // 8)
                    off = self._uint(body, p, nb)
                    p += nb
                dims = ()
                if ver == 1:
                    ndim = body[p]
                    p += 12
                    dims = tuple(self._uint(body, p + 4 * i, 4) for i in range(ndim))
                    p += 16
                mdt, used = self._parse_datatype(body, p)
                p += used
                if dims:
                    mdt = np.dtype((mdt, dims))
                names.append(name)
                formats.append(mdt)
                offsets.append(off)
            dt = np.dtype({"names": names, "formats": formats, "offsets": offsets,
                           "itemsize": size})
            return dt, p - pos
        if cls == 10:
            rank = body[p]
            p += 4 if ver < 3 else 1
            dims = tuple(self._uint(body, p + 4 * i, 4) for i in range(rank))
            p += 4 * rank
            if ver < 3:
                p += 4 * rank  # permutation indices
            base, used = self._parse_datatype(body, p)
            p += used
            return np.dtype((base, dims)), p - pos
        raise H5FormatError(f"datatype class {cls} not supported")

    def _parse_layout(self, body):
        ver = body[0]
        if ver == 3:
            cls = body[1]
            if cls == 0:
                size = self._uint(body, 2, 2)
                return ("compact", bytes(body[4:4 + size]))
            if cls == 1:
                addr = self._uint(body, 2, self._so)
                size = self._uint(body, 2 + self._so, self._sl)
                return ("contiguous", addr, size)
            if cls == 2:
                ndim = body[2]
                addr = self._uint(body, 3, self._so)
                dims = tuple(self._uint(body, 3 + self._so + 4 * i, 4) for i in range(ndim))
                return ("chunked", addr, dims)
        elif ver in (1, 2):
            ndim, cls = body[1], body[2]
            pos = 8
            addr = None
            if cls != 0:
                addr = self._uint(body, pos, self._so)
                pos += self._so
            dims = tuple(self._uint(body, pos + 4 * i, 4) for i in range(ndim))
            pos += 4 * ndim
            if cls == 1:
                return ("contiguous", addr, None)
            if cls == 2:
                return ("chunked", addr, dims)
            size = self._uint(body, pos, 4)
            return ("compact", bytes(body[pos + 4:pos + 4 + size]))
        raise H5FormatError(f"data layout version {ver} not supported")

    def _parse_filters(self, body):
        ver, nfilt = body[0], body[1]
        pos = 8 if ver == 1 else 2
        out = []
        for _ in range(nfilt):
            fid = self._uint(body, pos, 2)
            if ver == 1 or fid >= 256:
                nlen = self._uint(body, pos + 2, 2)
                pos += 4
            else:
                nlen = 0
                pos += 2
            _flags = self._uint(body, pos, 2)
            ncd = self._uint(body, pos + 2, 2)
            pos += 4
            if ver == 1:
                nlen = (nlen + 7) // 8 * 8
            pos += nlen
            cd = [self._uint(body, pos + 4 * i, 4) for i in range(ncd)]
            pos += 4 * ncd
            if ver == 1 and ncd % 2:
                pos += 4
            out.append((fid, cd))
        return out

    def _parse_attribute(self, body):
        ver = body[0]
        nsize = self._uint(body, 2, 2)
        tsize = self._uint(body, 4, 2)
        ssize = self._uint(body, 6, 2)
        pos = 8 + (1 if ver == 3 else 0)

        def pad(n):
            return (n + 7) // 8 * 8 if ver == 1 else n

        name = body[pos:pos + nsize].split(b"\0")[0].decode("utf8")
        pos += pad(nsize)
        tbody = body[pos:pos + tsize]
        pos += pad(tsize)
        sbody = body[pos:pos + ssize]
        pos += pad(ssize)
        if (tbody[0] & 0x0F) == 9:
            return name, None  # variable-length attribute: not needed by the path
        dt, _ = self._parse_datatype(tbody, 0)
        shape = self._parse_dataspace(sbody) if sbody[1] else ()
        n = int(np.prod(shape)) if shape else 1
        val = np.frombuffer(body, dtype=dt, count=n, offset=pos)
        if dt.kind == "S":
            val = [v.split(b"\0")[0].decode("utf8") for v in val]
            return name, (val[0] if not shape else val)
        return name, (val[0] if not shape else val.copy().reshape(shape))

    # -------------------------------------------------------------- chunked read
    def _read_chunked(self, ds, rows=None):
        _, btree, cdims = ds._layout
        rank = len(ds.shape)
        chunk_shape = cdims[:rank]
        r0, r1 = (0, ds.shape[0]) if rows is None else rows
        out = np.zeros((r1 - r0,) + tuple(ds.shape[1:]), dtype=ds.dtype)
        if btree == _UNDEF:
            return out
        chunks = []
        self._walk_chunk_btree(btree, rank, chunks)
        itemsize = ds.dtype.itemsize
        for size, mask, offs, addr in chunks:
            if offs[0] >= r1 or offs[0] + chunk_shape[0] <= r0:
                continue                             # no row of the requested range in this chunk
            raw = self._buf[addr:addr + size]
            for idx, (fid, cd) in reversed(list(enumerate(ds._filters))):
                if mask & (1 << idx):
                    continue
                if fid == 1:
                    raw = zlib.decompress(raw)
                elif fid == 2:
                    a = np.frombuffer(raw, dtype=np.uint8)
                    es = cd[0] if cd else itemsize
                    nel = len(a) // es
                    raw = a[:nel * es].reshape(es, nel).T.tobytes() + a[nel * es:].tobytes()
                elif fid == 3:
                    raw = raw[:-4]  # fletcher32 checksum trailer
                else:
                    raise H5FormatError(f"filter {fid} not supported")
            block = np.frombuffer(raw, dtype=ds.dtype,
                                  count=int(np.prod(chunk_shape))).reshape(chunk_shape)
            sl_out, sl_in = [], []
            for axis, (o, c, n) in enumerate(zip(offs, chunk_shape, ds.shape)):
                lo, hi = o, min(o + c, n)
                if axis == 0:
                    lo, hi = max(lo, r0), min(hi, r1)
                    sl_out.append(slice(lo - r0, hi - r0))
                else:
                    sl_out.append(slice(lo, hi))
                sl_in.append(slice(lo - o, hi - o))
            out[tuple(sl_out)] = block[tuple(sl_in)]
        return out

    def _walk_chunk_btree(self, addr, rank, out):
        buf = self._buf
        if buf[addr:addr + 4] != b"TREE":
            raise H5FormatError("bad chunk B-tree node")
        level = buf[addr + 5]
        used = self._uint(buf, addr + 6, 2)
        pos = addr + 8 + 2 * self._so
        keysize = 8 + 8 * (rank + 1)
        for _ in range(used):
            size, mask = struct.unpack_from("<II", buf, pos)
            offs = struct.unpack_from(f"<{rank}Q", buf, pos + 8)
            child = self._uint(buf, pos + keysize, self._so)
            pos += keysize + self._so
            if level > 0:
                self._walk_chunk_btree(child, rank, out)
            else:
                out.append((size, mask, offs, child))

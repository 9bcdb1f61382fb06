"""
Minimal HDF5 *appender* -- just enough to add solution tables to an existing h5parm.

The reference's KL fit leaves its result in the input file (stationscreen.py:1087-1159 through
``Solset.make_soltab``, h5parm.py:510-640): soltabs ``phase_screen000`` (values, weights after
flagging) and ``phase_screen000resid`` (residuals, the screen ORDER stored as weights), the
attributes ``beta / r_0 / height / midra / middec`` and the array ``piercepoint`` -- LoSoTo and the
reference's own ``KLScreen.fit`` read them back (kl_screen.py:113-154).  There is no HDF5 library
on this image, so the side effect is produced by this writer of the on-disk format (the same
subset ``h5min`` reads, which is what pytables / DP3 write for h5parms):

* superblock version 0/1 with 8-byte offsets and lengths
* "old style" groups: symbol-table message -> v1 B-tree (``TREE``) -> ``SNOD`` -> local heap
* version-1 object headers; contiguous datasets; v1 dataspaces; IEEE float (f16 / f32 / f64),
  fixed-point and fixed-length string datatypes; v1 attribute messages

How a group gains a child without a library: every new object (raw data, object headers, the
child group's own heap / SNOD / B-tree) is appended at the end of the file; then the PARENT
group's symbol table is rebuilt there too -- a fresh local heap, symbol node(s) and B-tree root
holding the old entries plus the new one -- and three fixed-size fields are patched in place:
the parent's symbol-table message (B-tree and heap address), the cached copy of those two
addresses in the grandparent's symbol-table entry, and the superblock's end-of-file address.
The old table becomes dead space.  New bytes are written and flushed before any pointer moves,
so an interrupted call leaves the old, valid file.

An existing child of the same name is REPLACED (the reference's ``remove_soltabs`` never
succeeds and ``make_soltab`` then writes under another name while ``get_soltab`` keeps reading
the stale table; not reproduced).
"""
import os
import struct

import numpy as np

from . import h5min

_UNDEF = 0xFFFFFFFFFFFFFFFF


def _pad8(data):
    return data + b"\0" * (-len(data) % 8)


# ------------------------------------------------------------------------------ messages
def _datatype(dt):
    """Datatype message body (version 1) for a numpy dtype."""
    dt = np.dtype(dt)
    if dt.kind == "f":
        layout = {2: (15, 10, 5, 10, 15), 4: (31, 23, 8, 23, 127), 8: (63, 52, 11, 52, 1023)}[dt.itemsize]
        sign, eloc, esize, msize, bias = layout
        # class 1 (floating point); bit field: little-endian, mantissa normalisation 2 (implied msb)
        head = bytes([0x11, 0x20, sign, 0x00]) + struct.pack("<I", dt.itemsize)
        return head + struct.pack("<HHBBBBI", 0, dt.itemsize * 8, eloc, esize, 0, msize, bias)
    if dt.kind in "iu":
        head = bytes([0x10, 0x08 if dt.kind == "i" else 0x00, 0x00, 0x00]) + struct.pack("<I", dt.itemsize)
        return head + struct.pack("<HH", 0, dt.itemsize * 8)
    if dt.kind == "S":
        # class 3 (string): null-padded (1), ASCII
        return bytes([0x13, 0x01, 0x00, 0x00]) + struct.pack("<I", dt.itemsize)
    raise TypeError(f"datatype {dt} not supported by the h5parm appender")


def _dataspace(shape):
    """Dataspace message body, version 1, no maximum dimensions."""
    return bytes([1, len(shape), 0, 0, 0, 0, 0, 0]) + b"".join(struct.pack("<Q", int(n)) for n in shape)


def _as_array(value):
    if isinstance(value, str):
        value = value.encode("utf8")
    if isinstance(value, bytes):
        return np.array(value, dtype=f"S{max(1, len(value))}")
    arr = np.asarray(value)
    if arr.dtype.kind == "U":
        arr = np.char.encode(arr, "utf8")
    if arr.dtype.kind == "f" and arr.dtype.itemsize not in (2, 4, 8):
        arr = arr.astype(np.float64)
    if arr.dtype.kind == "b":
        arr = arr.astype(np.int8)
    return arr.astype(arr.dtype.newbyteorder("<"), copy=False)


def _attribute(name, value):
    """Attribute message body, version 1."""
    arr = _as_array(value)
    nm = name.encode("utf8") + b"\0"
    dtb = _datatype(arr.dtype)
    dsb = _dataspace(arr.shape)
    head = struct.pack("<BBHHH", 1, 0, len(nm), len(dtb), len(dsb))
    return head + _pad8(nm) + _pad8(dtb) + _pad8(dsb) + np.ascontiguousarray(arr).tobytes()


def _object_header(messages):
    """Version-1 object header from [(type, body)] (bodies are padded to 8 bytes)."""
    body = b""
    for mtype, data in messages:
        data = _pad8(data)
        body += struct.pack("<HHB3x", mtype, len(data), 0) + data
    return struct.pack("<BBHII4x", 1, 0, len(messages), 1, len(body)) + body


# ------------------------------------------------------------------------------ appender
class Appender:
    """Appends objects to an existing HDF5 file of the h5parm subset (see the module text)."""

    def __init__(self, path):
        self.path = path
        with h5min.File(path) as handle:
            if handle._so != 8 or handle._sl != 8:
                raise h5min.H5FormatError("the appender needs 8-byte offsets and lengths")
            buf = handle._buf
            base = 0
            while buf[base:base + 8] != h5min._SIG:
                base = 512 if base == 0 else base * 2
            self._sb_version = buf[base + 8]
            self._leaf_k = int.from_bytes(buf[base + 16:base + 18], "little")
            self._internal_k = int.from_bytes(buf[base + 18:base + 20], "little")
            fields = base + 24 + (4 if self._sb_version == 1 else 0)
            if int.from_bytes(buf[fields:fields + 8], "little") != 0:
                raise h5min.H5FormatError("non-zero base address not supported")
            self._eof_field = fields + 16
            self._root_entry = fields + 32           # root group symbol-table entry
            self._eof = max(int.from_bytes(buf[self._eof_field:self._eof_field + 8], "little"), len(buf))
        self._fd = os.open(path, os.O_RDWR)
        self._eof = max(self._eof, os.fstat(self._fd).st_size)
        self._patches = []

    # -- raw placement
    def _put(self, data):
        addr = (self._eof + 7) & ~7
        os.pwrite(self._fd, data, addr)
        self._eof = addr + len(data)
        return addr

    def _dataset(self, array, attrs=None):
        arr = _as_array(array)
        raw = np.ascontiguousarray(arr).tobytes()
        addr = self._put(raw) if raw else _UNDEF
        msgs = [(0x01, _dataspace(arr.shape)), (0x03, _datatype(arr.dtype)),
                (0x05, bytes([2, 1, 0, 0])),                       # fill value v2: early allocation, undefined
                (0x08, bytes([3, 1]) + struct.pack("<QQ", addr, len(raw)))]
        msgs += [(0x0C, _attribute(k, v)) for k, v in (attrs or {}).items()]
        return self._put(_object_header(msgs))

    def _symbol_table(self, entries):
        """entries: {name: (object header address, (btree, heap) or None)} -> (btree, heap)."""
        names = sorted(entries, key=lambda s: s.encode("utf8"))
        data = bytearray(8)                                        # offset 0: the empty string
        offs = {}
        for name in names:
            offs[name] = len(data)
            data += _pad8(name.encode("utf8") + b"\0")
        data_addr = self._put(bytes(data))
        heap = self._put(b"HEAP" + bytes(4) + struct.pack("<QQQ", len(data), 1, data_addr))   # free list: none
        per = 2 * self._leaf_k
        groups = [names[i:i + per] for i in range(0, len(names), per)] or [[]]
        if len(groups) > 2 * self._internal_k:
            raise h5min.H5FormatError("too many children for one B-tree node")
        children = []
        for grp in groups:
            node = b"SNOD" + struct.pack("<BBH", 1, 0, len(grp))
            for name in grp:
                ohdr, stab = entries[name]
                if stab is None:
                    node += struct.pack("<QQII16x", offs[name], ohdr, 0, 0)
                else:
                    node += struct.pack("<QQIIQQ", offs[name], ohdr, 1, 0, stab[0], stab[1])
            node += bytes(40 * (per - len(grp)))
            children.append((self._put(node), offs[grp[-1]] if grp else 0))
        tree = b"TREE" + struct.pack("<BBHQQ", 0, 0, len(children), _UNDEF, _UNDEF) + struct.pack("<Q", 0)
        for addr, last_key in children:
            tree += struct.pack("<QQ", addr, last_key)
        tree += bytes(16 * (2 * self._internal_k - len(children)))
        return self._put(tree), heap

    def _group(self, title, attrs, children):
        """children: {name: ("dataset", array, attrs) | ("addr", ohdr, stab)}."""
        entries = {}
        for name, spec in children.items():
            if spec[0] == "dataset":
                entries[name] = (self._dataset(spec[1], spec[2]), None)
            else:
                entries[name] = (spec[1], spec[2])
        btree, heap = self._symbol_table(entries)
        all_attrs = {"CLASS": "GROUP", "VERSION": "1.0", "TITLE": title}
        all_attrs.update(attrs or {})
        msgs = [(0x11, struct.pack("<QQ", btree, heap))] + [(0x0C, _attribute(k, v)) for k, v in all_attrs.items()]
        return self._put(_object_header(msgs)), (btree, heap)

    # -- public
    def add_group(self, parent_path, name, title, attrs, datasets):
        """Adds (or replaces) group ``name`` under the existing group ``parent_path``.
        ``datasets``: {name: array} or {name: (array, {attr: value})}."""
        children = {}
        for key, spec in datasets.items():
            array, dattrs = spec if isinstance(spec, tuple) else (spec, None)
            children[key] = ("dataset", array, dattrs)
        ohdr, stab = self._group(title, attrs, children)
        with h5min.File(self.path) as handle:
            parent = handle[parent_path] if parent_path.strip("/") else handle
            if not isinstance(parent, h5min.Group):
                raise h5min.H5FormatError(f"{parent_path} is not a group")
            old = handle._symbol_entries(parent)
            msg_off = handle._message_offset(parent.addr, 0x11)
            parts = [p for p in parent_path.split("/") if p]
            if parts:
                grand = handle["/".join(parts[:-1])] if len(parts) > 1 else handle
                entry_off = handle._entry_offset(grand, parts[-1])
            else:
                entry_off = self._root_entry
        entries = {k: (v[0], v[1]) for k, v in old.items()}
        entries[name] = (ohdr, stab)
        btree, heap = self._symbol_table(entries)
        os.fsync(self._fd)                                         # new bytes first, pointers after
        self._patch(msg_off, struct.pack("<QQ", btree, heap))
        self._patch(entry_off + 16, struct.pack("<IIQQ", 1, 0, btree, heap))
        self._patch(self._eof_field, struct.pack("<Q", (self._eof + 7) & ~7))
        os.ftruncate(self._fd, (self._eof + 7) & ~7)
        os.fsync(self._fd)

    def _patch(self, offset, data):
        os.pwrite(self._fd, data, offset)

    def close(self):
        if self._fd is not None:
            os.close(self._fd)
            self._fd = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False


def append_soltab(path, solset, name, soltype, axes_names, axes_vals, val, weight, attrs=None, extra=None,
                  weight_dtype=np.float16):
    """``Solset.make_soltab`` (h5parm.py:510-640) on a file: group ``/{solset}/{name}`` titled
    ``soltype`` with the axis arrays, ``val`` (float64) and ``weight`` (float16 like the
    reference's default), both tagged ``AXES``; ``attrs`` become attributes of the group,
    ``extra`` further arrays in it (the ``piercepoint`` table, stationscreen.py:1148-1154)."""
    val = np.asarray(val, np.float64)
    weight = np.asarray(weight).astype(weight_dtype)
    if list(val.shape) != [len(v) for v in axes_vals] or weight.shape != val.shape:
        raise ValueError("val / weight do not match the axes")
    axes = ",".join(axes_names)
    datasets = {n: np.asarray(v) for n, v in zip(axes_names, axes_vals)}
    datasets["val"] = (val, {"AXES": axes})
    datasets["weight"] = (weight, {"AXES": axes})
    for key, array in (extra or {}).items():
        datasets[key] = np.asarray(array)
    group_attrs = {"parmdb_type": ""}
    group_attrs.update(attrs or {})
    with Appender(path) as app:
        app.add_group(solset, name, soltype, group_attrs, datasets)

"""
TEST INFRASTRUCTURE (oracle): in-memory, duck-typed replacement of the reference's
``utils/h5parm.py`` (LoSoTo object model over pytables -- pytables is not installed).

Only the members the hot path touches exist (see fake_misc.py header for the call sites).
Reads go through the product's format parser ``ska_sdp_screen_fitting_b200.utils.h5min``
(pure I/O, pinned by the known answers of SURVEY.md appendix B); soltabs created by the
fit (``make_soltab``) live in memory, keyed by file name, for the life of the process.
"""
import numpy as np

from ska_sdp_screen_fitting_b200.utils import h5min

_VIRTUAL = {}  # filename -> {solset: dict(ant={name: xyz}, source={name: radec})}: h5parms that
               # exist only in memory (synthetic inputs; their soltabs are added with add_soltab)
_CREATED = {}  # filename -> {solset: {soltab_name: _Obj}}
_EXTRA = {}    # filename -> {solset: {soltab_name: dict(type, axes, axes_vals, val, weight)}}:
               # in-memory soltabs overlaid on a real file (synthetic amplitude tables)


def add_virtual(filename, solset, ant, source):
    """Declare an in-memory h5parm ``filename`` with one solset: ``ant`` = {name: float32 xyz},
    ``source`` = {name: float32 [ra, dec] rad} (the antenna / source tables, h5parm.py:731-767)."""
    _VIRTUAL.setdefault(filename, {})[solset] = dict(ant=dict(ant), source=dict(source))


def add_soltab(filename, solset, name, soltype, axes, axes_vals, val, weight):
    """Overlay an in-memory soltab on ``filename`` (there is no HDF5 writer in this image)."""
    _EXTRA.setdefault(filename, {}).setdefault(solset, {})[name] = dict(
        type=soltype, axes=list(axes), axes_vals=[np.asarray(v) for v in axes_vals],
        val=np.asarray(val), weight=np.asarray(weight))


class _VFile:
    def __init__(self, store):
        self._store = store

    def create_array(self, where, name, obj=None):
        parts = [p for p in where.split("/") if p]
        node = self._store[parts[0]][parts[1]]
        setattr(node, name, np.array(obj))


class _Obj:
    """Stands in for the pytables Group of one soltab."""

    def __init__(self, name, title, store):
        self._v_name = name
        self._v_title = title
        self._v_attrs = {}
        self._v_file = _VFile(store)


class Soltab:
    def __init__(self, solset, name, soltype, axes_names, axes_vals, vals, weights, obj):
        self._solset = solset
        self.name = name
        self._type = soltype
        self.axes_names = list(axes_names)
        self._axes = {n: np.asarray(v) for n, v in zip(axes_names, axes_vals)}
        self._val = np.asarray(vals)
        self._weight = np.asarray(weights)
        self.obj = obj

    def __getattr__(self, axis):
        # mirrors h5parm.py:1306-1322
        if axis == "val":
            return self.get_values(ret_axes_vals=False)
        if axis == "weight":
            return self.get_values(ret_axes_vals=False, weight=True)
        if axis in self.__dict__.get("axes_names", ()):
            return self._axes[axis].copy()
        raise AttributeError(axis)

    def get_values(self, ret_axes_vals=True, weight=False):
        data = (self._weight if weight else self._val).copy()
        if ret_axes_vals:
            return data, {k: v.copy() for k, v in self._axes.items()}
        return data

    def get_type(self):
        return self._type

    def get_axes_names(self):
        return self.axes_names[:]

    def get_solset(self):
        return self._solset

    def add_history(self, entry, date=True):
        del entry, date


class Solset:
    def __init__(self, h5, name):
        self._h5 = h5
        self.name = name
        self.obj = _Obj(name, "", h5._store)

    def _file_group(self):
        return self._h5._file[self.name]

    def get_soltab(self, name, use_cache=False, sel=None):
        del use_cache, sel
        created = self._h5._store.setdefault(self.name, {})
        if name in created:
            return created[name]._soltab
        extra = _EXTRA.get(self._h5.filename, {}).get(self.name, {})
        if name in extra:
            e = extra[name]
            obj = _Obj(name, e["type"], self._h5._store)
            return Soltab(self, name, e["type"], e["axes"], e["axes_vals"], e["val"].copy(),
                          e["weight"].copy(), obj)
        if self._h5._file is None:
            raise Exception(f"Solution-table {name}  not found in solset {self.name}.")
        grp = self._file_group()
        if name not in grp:
            raise Exception(f"Solution-table {name}  not found in solset {self.name}.")
        tab = grp[name]
        axes = tab["val"].attrs["AXES"].split(",")
        vals = []
        for ax in axes:
            a = tab[ax].read()
            if a.dtype.kind == "S":
                a = np.array([s.split(b"\0")[0].decode() for s in a])
            vals.append(a)
        obj = _Obj(name, tab.attrs.get("TITLE", ""), self._h5._store)
        return Soltab(self, name, tab.attrs.get("TITLE", ""), axes, vals,
                      tab["val"].read(), tab["weight"].read(), obj)

    def get_soltab_names(self):
        if self._h5._file is None:
            names = list(_EXTRA.get(self._h5.filename, {}).get(self.name, {}).keys())
        else:
            names = [k for k in self._file_group().keys() if k not in ("antenna", "source")]
        return names + list(self._h5._store.get(self.name, {}).keys())

    def get_ant(self):
        if self._h5._file is None:
            return dict(_VIRTUAL[self._h5.filename][self.name]["ant"])
        out = {}
        for row in self._file_group()["antenna"].read():
            out[row["name"].split(b"\0")[0].decode()] = row["position"]
        return out

    def get_source(self):
        if self._h5._file is None:
            return dict(_VIRTUAL[self._h5.filename][self.name]["source"])
        out = {}
        for row in self._file_group()["source"].read():
            out[row["name"].split(b"\0")[0].decode()] = row["dir"]
        return out

    def make_soltab(self, soltype, soltab_name, axes_names, axes_vals, vals, weights):
        created = self._h5._store.setdefault(self.name, {})
        obj = _Obj(soltab_name, soltype, self._h5._store)
        st = Soltab(self, soltab_name, soltype, axes_names, axes_vals,
                    np.array(vals), np.array(weights), obj)
        obj._soltab = st
        created[soltab_name] = obj
        return st


class H5parm:
    def __init__(self, h5parm_file, readonly=True, complevel=0, complib="zlib"):
        del readonly, complevel, complib
        self.filename = h5parm_file
        self._file = None if h5parm_file in _VIRTUAL else h5min.File(h5parm_file)
        self._store = _CREATED.setdefault(h5parm_file, {})

    def get_solset(self, name):
        if name not in (self._file if self._file is not None else _VIRTUAL[self.filename]):
            raise Exception(f"Solution set {name} not found.")
        return Solset(self, name)

    def close(self):
        pass


def reset():
    _CREATED.clear()
    _EXTRA.clear()
    _VIRTUAL.clear()

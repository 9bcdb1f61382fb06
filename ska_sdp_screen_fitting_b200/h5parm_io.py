"""
Read contract of the reference's ``utils/h5parm.py`` that the hot path needs
(h5parm.py:107-120 open, :731-767 antenna / source tables, :1306-1322 axis datasets,
:1413-1762 values in AXES order), served by the dependency-free HDF5 reader ``utils.h5min``.

``Solutions`` is a plain in-memory record; it can also be built directly from arrays
(``Solutions.from_arrays``) which is how synthetic benchmark inputs enter the engine
without an HDF5 writer.
"""
from dataclasses import dataclass, field

import numpy as np

from .utils import h5min


def _text(values):
    out = []
    for v in np.asarray(values).ravel():
        if isinstance(v, bytes):
            v = v.decode("utf8", "replace")
        out.append(str(v).rstrip("\x00 "))
    return out


@dataclass
class Solutions:
    """One soltab with the tables of its solset.  ``val``/``weight`` are in canonical axis
    order [time, freq, ant, dir(, pol)]."""
    soltype: str
    val: np.ndarray
    weight: np.ndarray
    time: np.ndarray
    freq: np.ndarray
    ant: list
    dir: list
    pol: list = field(default_factory=list)
    source_dict: dict = field(default_factory=dict)      # name -> float32 [ra, dec] rad
    station_dict: dict = field(default_factory=dict)     # name -> float32 [x, y, z] m

    @property
    def source_positions(self):
        return [self.source_dict[name] for name in self.dir]

    @property
    def station_positions(self):
        return [self.station_dict[name] for name in self.ant]

    def rows(self, time_ranges):
        """The same soltab restricted to the given (start, stop) time-index ranges (views when
        there is one range)."""
        if len(time_ranges) == 1:
            a, b = time_ranges[0]
            val, weight, time = self.val[a:b], self.weight[a:b], self.time[a:b]
        else:
            idx = np.concatenate([np.arange(a, b) for a, b in time_ranges]) if time_ranges else np.arange(0)
            val, weight, time = self.val[idx], self.weight[idx], self.time[idx]
        return Solutions(self.soltype, val, weight, time, self.freq, self.ant, self.dir, self.pol,
                         self.source_dict, self.station_dict)

    @classmethod
    def from_arrays(cls, val, weight, time, freq, ant, dirs, source_rad, station_xyz,
                    soltype="phase"):
        ant, dirs = list(ant), list(dirs)
        return cls(soltype, np.asarray(val), np.asarray(weight), np.asarray(time, np.float64),
                   np.asarray(freq, np.float64), ant, dirs, [],
                   {n: np.asarray(p, np.float32) for n, p in zip(dirs, source_rad)},
                   {n: np.asarray(p, np.float32) for n, p in zip(ant, station_xyz)})


def count_directions(filename, solset_name, soltab_name):
    """make_aterm_images.py:107-113."""
    with h5min.File(filename) as handle:
        return len(np.asarray(handle[f"{solset_name}/{soltab_name}/dir"]))


def read_time_axis(filename, solset_name="sol000", soltab_name="phase000"):
    """The (small) time axis alone: every rank of a time-sharded run reads it in full to agree on
    the output-file boundaries (screen.py:283-294) before reading its own rows of val/weight."""
    with h5min.File(filename) as handle:
        return np.asarray(handle[f"{solset_name}/{soltab_name}/time"].read(), np.float64)


def read_solutions(filename, solset_name="sol000", soltab_name="phase000", time_ranges=None):
    """``time_ranges``: optional list of (start, stop) time-index ranges; only those rows of
    val / weight are read (each rank reads only its time rows) and ``time`` is
    restricted to match."""
    with h5min.File(filename) as handle:
        if solset_name not in handle:
            raise KeyError(f"Solution set {solset_name} not found in {filename}")
        solset = handle[solset_name]
        if soltab_name not in solset:
            raise KeyError(f"Solution table {soltab_name} not found in {solset_name}")
        tab = solset[soltab_name]
        val_ds, wt_ds = tab["val"], tab["weight"]
        axes = val_ds.attrs.get("AXES")
        if isinstance(axes, bytes):
            axes = axes.decode()
        axes = [a.strip() for a in str(axes).split(",")]
        want = [a for a in ("time", "freq", "ant", "dir", "pol") if a in axes]
        missing = [a for a in ("time", "freq", "ant", "dir") if a not in axes]
        if missing:
            raise ValueError(f"{soltab_name}: axes {missing} missing (have {axes})")
        perm = [axes.index(a) for a in want]
        times = np.asarray(tab["time"].read(), np.float64)
        if time_ranges is None:
            val = np.asarray(val_ds.read()).transpose(perm)
            weight = np.asarray(wt_ds.read()).transpose(perm)
        elif axes[0] == "time":
            # time-major on disk (what DP3 and LoSoTo write): read just the rows
            val = np.concatenate([val_ds.read_rows(a, b) for a, b in time_ranges]).transpose(perm)
            weight = np.concatenate([wt_ds.read_rows(a, b) for a, b in time_ranges]).transpose(perm)
        else:
            rows = np.concatenate([np.arange(a, b) for a, b in time_ranges])
            val = np.asarray(val_ds.read()).transpose(perm)[rows]
            weight = np.asarray(wt_ds.read()).transpose(perm)[rows]
        if time_ranges is not None:
            times = np.concatenate([times[a:b] for a, b in time_ranges]) if time_ranges else times[:0]
        title = tab.attrs.get("TITLE", "phase")
        if isinstance(title, bytes):
            title = title.decode()
        source = np.asarray(solset["source"].read()) if "source" in solset else []
        antenna = np.asarray(solset["antenna"].read()) if "antenna" in solset else []
        return Solutions(
            soltype=str(title),
            val=np.ascontiguousarray(val),
            weight=np.ascontiguousarray(weight),
            time=times,
            freq=np.asarray(tab["freq"].read(), np.float64),
            ant=_text(tab["ant"].read()),
            dir=_text(tab["dir"].read()),
            pol=_text(tab["pol"].read()) if "pol" in tab else [],
            source_dict={n: np.asarray(r["dir"], np.float32) for n, r in
                         zip(_text([r["name"] for r in source]), source)},
            station_dict={n: np.asarray(r["position"], np.float32) for n, r in
                          zip(_text([r["name"] for r in antenna]), antenna)},
        )

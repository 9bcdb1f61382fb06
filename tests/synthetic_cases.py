"""
TEST / BENCHMARK INFRASTRUCTURE (not part of the product package).

Seeded synthetic solution sets in the shapes BASELINE.json names: LOFAR-HBA
(62 stations, 45 directions), SKA-Low (512 stations, 100 directions) and small test shapes.
Everything is generated from ``numpy.random.default_rng(seed)`` and recorded in the result.
"""
from dataclasses import dataclass

import numpy as np

from ska_sdp_screen_fitting_b200.h5parm_io import Solutions

FIELD_RA, FIELD_DEC = 126.23, 64.50
T0, DT = 4.98795843e9, 8.0111208
F0, DF = 122.215271e6, 4.0527e6
CS001 = np.array([3826896.2, 460979.5, 5064658.0])

# name -> (stations, directions, times, freqs, image size, cell size deg)
SHAPES = {
    "tiny": (6, 5, 4, 2, 16, 0.2),
    "small": (12, 9, 8, 3, 64, 0.05),
    "lofar": (62, 45, 720, 24, 512, 0.01),        # BASELINE.json configs[2] (C3)
    "voronoi1k": (62, 45, 720, 24, 1024, 0.02),   # configs[3] (C4)
    "skalow": (512, 100, 2048, 32, 2048, 0.005),  # configs[4] (C5)
}


@dataclass
class Case:
    name: str
    seed: int
    solutions: Solutions
    patch_radec: np.ndarray      # [ndir, 2] deg (the "sky model" positions, float64)
    rad: float
    dec: float
    width_deg: float
    cellsize_deg: float
    n_pix: int


def make_case(name="small", seed=3, n_times=None, n_freqs=None, flag_fraction=0.01, n_pix=None,
              cellsize_deg=None, n_stations=None, n_dirs=None, time_offset=0):
    """Builds a synthetic case.  ``n_times`` / ``n_freqs`` / ... override the table (benchmarks
    take a time slab of the full configuration, ``time_offset`` = index of its first slot).  Width is (n - 0.5) * cell so that
    ceil(width / cell) == n exactly."""
    st, nd, nt, nf, npx, cell = SHAPES[name]
    st = n_stations or st
    nd = n_dirs or nd
    nt = n_times or nt
    nf = n_freqs or nf
    npx = n_pix or npx
    cell = cellsize_deg or cell
    rng = np.random.default_rng(seed)
    width = (npx - 0.5) * cell
    times = T0 + DT * (time_offset + np.arange(nt))
    freqs = F0 + DF * np.arange(nf)
    n_core = min(48, st)
    xyz = np.vstack([CS001 + rng.normal(0.0, 3e3, (n_core, 3)),
                     CS001 + rng.normal(0.0, 4e4, (st - n_core, 3))])
    xyz[0] = CS001
    # directions uniform in the central 80 % of the field
    off_dec = rng.uniform(-0.4, 0.4, nd) * width
    off_ra = rng.uniform(-0.4, 0.4, nd) * width / np.cos(np.deg2rad(FIELD_DEC))
    radec = np.column_stack([FIELD_RA + off_ra, FIELD_DEC + off_dec])
    source_rad = np.deg2rad(radec).astype(np.float32)
    xi = np.deg2rad(off_ra) * np.cos(np.deg2rad(FIELD_DEC)) * 50.0
    eta = np.deg2rad(off_dec) * 50.0
    dist = np.linalg.norm(xyz - xyz[0], axis=1)
    gain = dist / max(dist.max(), 1.0)
    coef = rng.normal(0.0, 1.0, 3)
    spatial = coef[0] * xi + coef[1] * eta + coef[2] * xi * eta                     # [d]
    # geometry above depends on `seed` only; the per-sample noise and flags also on the time
    # offset, so ranks that own different time slabs share one array / field geometry
    rng = np.random.default_rng([seed, time_offset])
    val = (gain[None, None, :, None] * spatial[None, None, None, :]
           * (freqs[0] / freqs)[None, :, None, None]
           + 0.3 * np.sin(2 * np.pi * (times - times[0])[:, None, None, None] / 600.0
                          + np.arange(st)[None, None, :, None])
           * gain[None, None, :, None]
           + rng.normal(0.0, 0.02, (nt, nf, st, nd)))
    weight = np.ones((nt, nf, st, nd), np.float32)
    if flag_fraction > 0:
        weight[rng.random((nt, nf, st, nd)) < flag_fraction] = 0.0
    ants = [f"ST{i:03d}" for i in range(st)]
    dirs = [f"[Patch_{i}]" for i in range(nd)]
    sol = Solutions.from_arrays(val, weight, times, freqs, ants, dirs, source_rad,
                                xyz.astype(np.float32))
    return Case(name, seed, sol, radec, FIELD_RA, FIELD_DEC, width, cell, npx)

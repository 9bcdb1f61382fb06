"""
TEST INFRASTRUCTURE: one rank of the world_size-2 ``gloo`` test (tests/test_distributed.py).

Runs the host-side sharding logic of ``Screen.write`` -- every rank owns a contiguous range of
the time slots of each output file and writes its own byte range of the one FITS cube -- with
the kernels supplied by the CPU emulator build of the kernel sources (no GPU in the build
container).  Usage: python dist_worker.py <rank> <world> <port> <emu_lib> <golden_dir> <out_dir>
"""
import os
import sys

import numpy as np


def main():
    rank, world, port = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
    emu_lib, golden, out_dir = sys.argv[4], sys.argv[5], sys.argv[6]
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests", "emu"))
    import emu_engine
    emu_engine.install(emu_lib)
    import torch
    import torch.distributed as dist

    from ska_sdp_screen_fitting_b200 import skymodel
    from ska_sdp_screen_fitting_b200.h5parm_io import Solutions, read_solutions
    from ska_sdp_screen_fitting_b200.kl_screen import KLScreen
    from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen

    if world > 1:
        dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    full = read_solutions(os.path.join(golden, "solutions.h5"), "sol000", "phase000")
    # a small cut of the repo data: 5 time slots (odd: uneven split over 2 ranks), 2 freqs, 5 stations
    t, f, a = slice(0, 5), slice(0, 2), slice(0, 5)
    times = np.array(full.time[t], float)
    times[3:] += 100.0                                  # a time gap -> two output files (3 + 2 slots)
    sol = Solutions.from_arrays(np.asarray(full.val)[t, f, a], np.asarray(full.weight)[t, f, a], times,
                                np.asarray(full.freq)[f], list(full.ant[a]), list(full.dir),
                                np.asarray(full.source_positions), np.asarray(full.station_positions)[a])
    positions = skymodel.patch_positions(os.path.join(golden, "skymodel.txt"))
    radec = np.array([positions[name.strip("[]")] for name in sol.dir])
    for cls, name, smooth in ((VoronoiScreen, "vor", 0.5), (KLScreen, "kl", 0.0)):
        scr = cls(name, None, None, 126.23, 64.50, 3.33, 3.33)
        scr.solutions = sol
        if cls is VoronoiScreen:
            scr.patch_radec = radec
        scr.process()
        if world > 1:
            # 3 columns per device slab: slabs end in the middle of a time slot (10 columns each)
            scr.slab_bytes = 3 * 16 * 17 * 17
        files = scr.write(out_dir, 0.2, smooth_pix=smooth)
        assert len(files) == 2
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

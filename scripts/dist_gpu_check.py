"""Multi-GPU check of the public write path on real devices (run under torchrun on N GPUs):
every rank fits on its own GPU, owns a contiguous range of the time slots of each output file
and writes its byte range; NCCL carries only barriers.  Rank 0 then rewrites the same cubes
alone and compares the files byte for byte.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29533 scripts/dist_gpu_check.py <out_dir>
"""
import filecmp
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

sys.path.insert(0, os.path.join(ROOT, "tests"))
import synthetic_cases as synthetic  # noqa: E402
from ska_sdp_screen_fitting_b200.engine import Engine  # noqa: E402
from ska_sdp_screen_fitting_b200.kl_screen import KLScreen  # noqa: E402
from ska_sdp_screen_fitting_b200.voronoi_screen import VoronoiScreen  # noqa: E402


def build(cls, name, case, eng, slab_bytes):
    scr = cls(name, None, None, case.rad, case.dec, case.width_deg, case.width_deg)
    scr._engine = eng
    scr.solutions = case.solutions
    scr.slab_bytes = slab_bytes
    if cls is VoronoiScreen:
        scr.patch_radec = case.patch_radec
        scr.template_mode = "nearest"
    scr.process()
    return scr


def main():
    out_dir = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine(device=f"cuda:{local}")
    # 11 time slots (uneven split), a gap after slot 6 -> two files; 128 x 128 images, 9 directions
    case = synthetic.make_case("small", seed=5, n_times=11, n_pix=128, cellsize_deg=0.025)
    case.solutions.time[7:] += 500.0
    many, one = os.path.join(out_dir, "many"), os.path.join(out_dir, "one")
    if rank == 0:
        os.makedirs(many, exist_ok=True)
        os.makedirs(one, exist_ok=True)
    dist.barrier()
    slab = 7 * 16 * 128 * 128                      # 7 columns per device slab: cuts inside a time slot
    for cls, name, smooth in ((KLScreen, "kl", 0.0), (VoronoiScreen, "vor", 1.5)):
        scr = build(cls, name, case, eng, slab)
        files = scr.write(many, case.cellsize_deg, smooth_pix=smooth)
        assert len(files) == 2, files
    dist.barrier()
    dist.destroy_process_group()                    # the single-rank rewrite must not see a process group
    if rank == 0:
        for cls, name, smooth in ((KLScreen, "kl", 0.0), (VoronoiScreen, "vor", 1.5)):
            scr = build(cls, name, case, eng, 2 << 30)
            scr.write(one, case.cellsize_deg, smooth_pix=smooth)
        names = sorted(n for n in os.listdir(one) if n.endswith(".fits"))
        assert names == sorted(n for n in os.listdir(many) if n.endswith(".fits")), (names, os.listdir(many))
        for n in names:
            assert filecmp.cmp(os.path.join(one, n), os.path.join(many, n), shallow=False), n
        print(f"OK: {world} ranks wrote {names} byte-identical to one rank "
              f"({sum(os.path.getsize(os.path.join(one, n)) for n in names) / 1e6:.1f} MB)")


if __name__ == "__main__":
    main()

"""
TEST INFRASTRUCTURE: build ``tests/emu/_build/libssf_emu.so`` -- the kernel sources of
``ska_sdp_screen_fitting_b200/csrc`` compiled by g++ against the CUDA emulator of
``cuda_emu.h`` -- so kernel LOGIC can be checked against the oracle in the GPU-less build
container.  Never shipped and never loaded by the product: tests hand its path to
``tests/emu/emu_engine.EmuEngine``.
"""
import hashlib
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
CSRC = os.path.join(ROOT, "ska_sdp_screen_fitting_b200", "csrc")
OUT = os.path.join(HERE, "_build")
LIB = os.path.join(OUT, "libssf_emu.so")
SOURCES = ["core.cu", "kldecomp.cu", "klfit.cu", "kleval.cu", "contract_simt.cu",
           "contract_tc.cu", "voronoi.cu", "smooth.cu"]
FLAGS = ["-O2", "-std=c++17", "-fPIC", "-x", "c++", "-include", os.path.join(HERE, "cuda_emu.h"),
         "-ffp-contract=off", "-Wno-unknown-pragmas", "-Wno-attributes"]


def _stamp():
    h = hashlib.sha256()
    for d, names in ((CSRC, sorted(os.listdir(CSRC))), (HERE, ["cuda_emu.h", "cuda_emu.cpp"])):
        for n in names:
            p = os.path.join(d, n)
            if os.path.isfile(p):
                with open(p, "rb") as fh:
                    h.update(n.encode() + fh.read())
    with open(os.path.join(ROOT, "include", "ssf.h"), "rb") as fh:
        h.update(fh.read())
    return h.hexdigest()


def build():
    os.makedirs(OUT, exist_ok=True)
    stamp_file = os.path.join(OUT, "stamp")
    stamp = _stamp()
    if os.path.exists(LIB) and os.path.exists(stamp_file):
        with open(stamp_file, encoding="utf8") as fh:
            if fh.read() == stamp:
                return LIB

    def cc(src):
        obj = os.path.join(OUT, src.replace(".cu", ".o"))
        res = subprocess.run(["g++", *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj],
                             capture_output=True, text=True)
        if res.returncode:
            raise RuntimeError(f"emu build failed on {src}:\n{res.stderr[-4000:]}")
        return obj

    with ThreadPoolExecutor(8) as pool:
        objs = list(pool.map(cc, SOURCES))
    rt = os.path.join(OUT, "cuda_emu.o")
    res = subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-c", os.path.join(HERE, "cuda_emu.cpp"),
                          "-o", rt], capture_output=True, text=True)
    if res.returncode:
        raise RuntimeError(res.stderr)
    res = subprocess.run(["g++", "-shared", "-o", LIB, *objs, rt], capture_output=True, text=True)
    if res.returncode:
        raise RuntimeError(res.stderr)
    with open(stamp_file, "w", encoding="utf8") as fh:
        fh.write(stamp)
    return LIB


if __name__ == "__main__":
    print(build())

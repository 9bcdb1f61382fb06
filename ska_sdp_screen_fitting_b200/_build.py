"""
In-tree build of ``libssf_b200.so`` (the C-ABI library of include/ssf.h) with nvcc for
sm_100a.  No torch headers are involved: the library is plain CUDA C++ with ``extern "C"``
entry points and is loaded through ctypes (``_lib.py``).

    python -m ska_sdp_screen_fitting_b200._build [--force]
"""
import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ_DIR = os.path.join(PKG, "csrc", "_obj")
LIB_PATH = os.path.join(PKG, "libssf_b200.so")

SOURCES = ["core.cu", "kldecomp.cu", "klfit.cu", "kleval.cu", "contract_simt.cu",
           "contract_tc.cu", "voronoi.cu", "smooth.cu"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
] + os.environ.get("SSF_NVCC_EXTRA", "").split()      # experiment hook (e.g. -DSSF_DECOMP_BIG=1024)


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libssf_b200.so can only be built with the CUDA toolkit")
    return exe


def _stamp():
    h = hashlib.sha256()
    names = sorted(os.listdir(CSRC))
    for name in names:
        if name.endswith((".cu", ".cuh", ".h")):
            with open(os.path.join(CSRC, name), "rb") as fh:
                h.update(name.encode())
                h.update(fh.read())
    with open(os.path.join(ROOT, "include", "ssf.h"), "rb") as fh:
        h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    """Compiles every .cu under csrc/ and links libssf_b200.so.  Returns the library path."""
    stamp_file = os.path.join(OBJ_DIR, "stamp")
    stamp = _stamp()
    if not force and os.path.exists(LIB_PATH) and os.path.exists(stamp_file):
        with open(stamp_file, encoding="utf8") as fh:
            if fh.read().strip() == stamp:
                return LIB_PATH
    os.makedirs(OBJ_DIR, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(res.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as pool:
        objs = list(pool.map(compile_one, SOURCES))
    cmd = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-cudart", "shared",
           "-Xlinker", "-rpath=/usr/local/cuda/lib64"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    with open(stamp_file, "w", encoding="utf8") as fh:
        fh.write(stamp)
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))

import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def emu_library():
    """libssf_emu.so: the product kernel sources compiled by g++ against the CUDA emulator
    (tests/emu).  TEST INFRASTRUCTURE -- lets kernel logic be checked without a GPU."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    import build_emu

    return build_emu.build()


@pytest.fixture(scope="session")
def emu_engine(emu_library):
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    from emu_engine import EmuEngine

    return EmuEngine(emu_library)


@pytest.fixture(scope="session")
def gpu_engine():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from ska_sdp_screen_fitting_b200 import _build
    from ska_sdp_screen_fitting_b200.engine import Engine

    if not os.path.exists(_build.LIB_PATH):
        _build.build()
    return Engine(device="cuda:0")


@pytest.fixture(scope="session")
def repo_solutions(golden_dir):
    from ska_sdp_screen_fitting_b200.h5parm_io import read_solutions

    return read_solutions(os.path.join(golden_dir, "solutions.h5"), "sol000", "phase000")

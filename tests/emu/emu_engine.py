"""
TEST INFRASTRUCTURE: an ``Engine`` whose kernels come from ``libssf_emu.so`` -- the product's
kernel sources compiled by g++ against the CUDA emulator (tests/emu) -- on torch's CPU device,
so kernel and host logic can be checked in the GPU-less build container.

The product package has no such door: ``Engine()`` loads ``libssf_b200.so`` and refuses any
device but CUDA.  Tests construct ``EmuEngine`` explicitly, or call ``install()`` to make every
``Screen`` created afterwards in this process use one (the CLI / multi-rank tests, which run the
public entry points in a subprocess: ``python tests/emu/emu_engine.py <lib> <module> args...``).
"""
import os
import runpy
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from ska_sdp_screen_fitting_b200 import _lib  # noqa: E402
from ska_sdp_screen_fitting_b200.engine import Engine  # noqa: E402


class EmuEngine(Engine):
    def __init__(self, library):  # pylint: disable=super-init-not-called
        self._init_state(_lib.load(library), torch.device("cpu"))


def install(library):
    """Every Screen of this process gets an EmuEngine (replaces the lazy ``Screen.engine``)."""
    from ska_sdp_screen_fitting_b200 import screen

    shared = EmuEngine(library)

    def engine(self):
        if self._engine is None:
            self._engine = shared
        return self._engine

    screen.Screen.engine = property(engine)
    return shared


if __name__ == "__main__":
    # python emu_engine.py <libssf_emu.so> <module> [args...]: run a module's __main__ on the emulator
    install(sys.argv[1])
    sys.argv = sys.argv[2:]
    runpy.run_module(sys.argv[0], run_name="__main__", alter_sys=True)

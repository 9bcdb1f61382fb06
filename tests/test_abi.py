"""The C-ABI library builds, loads, and exports exactly what include/ssf.h declares."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_functions():
    with open(os.path.join(ROOT, "include", "ssf.h"), encoding="utf8") as fh:
        text = re.sub(r"/\*.*?\*/", "", fh.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(ssf_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_hot_path():
    names = declared_functions()
    for must in ("ssf_kl_decompose", "ssf_kl_fit_pass", "ssf_kl_eval", "ssf_voronoi_assign",
                 "ssf_voronoi_fill", "ssf_gauss_smooth", "ssf_nan_fix", "ssf_kl_cube_host",
                 "ssf_voronoi_cube_host"):
        assert must in names


def test_product_library_builds_and_exports_every_symbol():
    from ska_sdp_screen_fitting_b200 import _build, _lib

    path = _build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in ssf.h but not exported"
    assert sorted(_lib.SIGNATURES) == declared_functions()
    lib.ssf_version.restype = ctypes.c_int
    assert lib.ssf_version() >= 100


def test_emulator_library_exports_every_symbol(emu_library):
    lib = ctypes.CDLL(emu_library)
    for name in declared_functions():
        assert hasattr(lib, name)


def test_missing_library_fails_loudly(tmp_path, monkeypatch):
    from ska_sdp_screen_fitting_b200 import _lib

    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.SsfLibraryError):
        _lib.load()


def test_no_cpu_path_in_the_product_package(monkeypatch):
    """Engine() takes CUDA devices only, no environment variable or argument selects another
    library, and the product package never imports the oracle."""
    import inspect

    from ska_sdp_screen_fitting_b200 import _lib, engine
    from ska_sdp_screen_fitting_b200.engine import Engine

    with pytest.raises(_lib.SsfLibraryError):
        Engine(device="cpu")
    assert list(inspect.signature(Engine.__init__).parameters) == ["self", "device"]
    monkeypatch.setenv("SSF_B200_LIBRARY", "/nonexistent/libssf_emu.so")
    assert _lib.load() is _lib.load(_lib.LIB_PATH)
    pkg = os.path.dirname(engine.__file__)
    for base, _, names in os.walk(pkg):
        for name in names:
            if name.endswith(".py"):
                with open(os.path.join(base, name), encoding="utf8") as fh:
                    text = fh.read()
                assert "SSF_B200_LIBRARY" not in text, name
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), name


def test_bad_arguments_return_status_not_crash(emu_engine):
    import torch

    from ska_sdp_screen_fitting_b200 import _lib

    pp = torch.zeros((200, 2), dtype=torch.float64)
    masks = torch.zeros((1, 2), dtype=torch.int64)
    with pytest.raises(_lib.SsfError) as err:
        emu_engine.kl_decompose(pp, masks)
    assert err.value.code == -1 and "ndir" in str(err.value)

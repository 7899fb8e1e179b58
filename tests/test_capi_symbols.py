"""CPU: the C-ABI library builds, loads, and exports every symbol include/oige_b200.h declares.
No compute call is made (there is no GPU here)."""
import ctypes
import os
import re

import helpers
from omniisaacgymenvs_b200 import _lib, build

HEADER = os.path.join(helpers.ROOT, "include", "oige_b200.h")


def _declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(oige_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_loads():
    path = build.build_library()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    lib.oige_abi_version.restype = ctypes.c_int
    assert lib.oige_abi_version() == _lib.ABI_VERSION == 4


def test_every_declared_symbol_is_exported():
    build.build_library()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared_functions()
    assert len(declared) >= 12
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared


def test_struct_layouts_match_header_sizes():
    # sizes computed by hand from the header: catches ctypes / C drift without running a kernel
    assert ctypes.sizeof(_lib.OigeTaskParams) == 8 * 4 + 16 * 4 + 4 * 16 + 4 * 32 * 4
    assert ctypes.sizeof(_lib.OigeState) == 7 * 8
    assert ctypes.sizeof(_lib.OigeTaskBuffers) == 12 * 8  # ABI 4: episode_return
    assert ctypes.sizeof(_lib.OigeResetInit) == 3 * 8
    assert ctypes.sizeof(_lib.OigeCartpoleParams) == 8 * 4
    assert ctypes.sizeof(_lib.OigeCartpoleBuffers) == 7 * 8
    assert ctypes.sizeof(_lib.OigeAnymalParams) == 4 * 4 + 22 * 4 + 3 * 12 * 4
    assert ctypes.sizeof(_lib.OigeAnymalBuffers) == 11 * 8
    L, K, S = _lib.MAX_LINKS, _lib.MAX_SPHERES, _lib.MAX_SENSORS
    expect = (16 + 3 * L * 4 + L * 4 * (3 + 4 + 3 + 1 + 3 + 6 + 6 + 3) + K * 4 * (1 + 3 + 1) + S * 4 + (3 + 8) * 4
              + (4 + 1) * 4 + 2 * _lib.MAX_PAIRS * 4 + L * 7 * 4)  # ABI 3: velocity limits, self-collision pairs / capsules
    assert ctypes.sizeof(_lib.OigeModelDesc) == expect


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    import pytest

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.load()


def test_host_tensor_is_rejected():
    import pytest
    import torch

    with pytest.raises(RuntimeError, match="CUDA tensor"):
        _lib.dptr(torch.zeros(4), torch.float32, "x")

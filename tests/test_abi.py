"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol the header declares
(no compute calls - there is no GPU here)."""
import ctypes
import os
import re

from conftest import ROOT
from ragen_b200 import _lib, build


def header_symbols():
    text = open(os.path.join(ROOT, "include", "ragen_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ragen_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_header_symbols():
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    syms = header_symbols()
    assert len(syms) >= 9
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/ragen_b200.h but not exported"


def test_binding_covers_header():
    assert sorted(_lib.SIGNATURES) == header_symbols()
    lib = _lib.load()
    assert lib.ragen_version() >= 100


def test_no_cpu_fallback_without_gpu():
    import pytest
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.RagenError):
        _lib.Context(0)


def test_sass_uses_tma_bulk_copies():
    """The step kernel stages tiles with 1-D TMA bulk copies: UBLKCP must appear in the SASS."""
    import shutil
    import subprocess

    if not shutil.which("cuobjdump"):
        import pytest
        pytest.skip("cuobjdump not available")
    sass = subprocess.run(["cuobjdump", "-sass", build.LIB_PATH], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass
    assert "sm_100a" in sass or "SM100" in sass.upper()

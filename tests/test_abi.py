"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/fs3.h declares, and creating a simulation without a GPU fails loudly (no CPU fallback)."""
import ctypes
import os
import re
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "fs3.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(fs3_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from fluidsimulation_b200 import build, LIB_PATH, ABI_SYMBOLS
    build.build()
    lib = ctypes.CDLL(LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "missing export: " + s
    assert sorted(ABI_SYMBOLS) == syms


def test_particle_layout_is_24_bytes():
    txt = open(os.path.join(ROOT, "include", "fs3.h")).read()
    assert "float X, Y, Z;" in txt and "float uX, uY, uZ;" in txt


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from fluidsimulation_b200 import FluidSim3, Fs3Error
    with pytest.raises(Fs3Error):
        FluidSim3(16, 16, 16, 16.0)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "fluidsimulation_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read().lower().replace("the oracle", ""), f

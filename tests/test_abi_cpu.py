"""CPU-side checks of the boundary: the C-ABI library builds, loads and exports every symbol include/nhstep.h declares,
and the product fails loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    g.build()
    import oceananigans_jl_b200 as oc
    return oc


def header_symbols():
    src = open(os.path.join(ROOT, "include", "nhstep.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nh_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    lib = ctypes.CDLL(built._lib.LIB_PATH)
    declared = header_symbols()
    assert len(declared) >= 35
    for s in declared:
        assert hasattr(lib, s), f"libnhstep.so does not export {s}"
    assert sorted(built._lib.SYMBOLS) == declared, "python binding and header disagree"


def test_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    assert built.device_count() == 0
    with pytest.raises(built.NhError, match="no CPU fallback"):
        built.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "oceananigans.jl_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in text and "from oracle" not in text, f


def test_host_mirror_argument_validation(built):
    oc = built
    with pytest.raises(ValueError):
        oc.WENO(order=4)
    with pytest.raises(ValueError):
        oc.Centered(order=3)
    with pytest.raises(ValueError):
        oc.RectilinearGrid(np.float64, size=(4, 4, 4), topology=("Periodic", "Periodic"))
    with pytest.raises(ValueError):
        oc.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1), x=(0, 1))

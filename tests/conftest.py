import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: longer CPU oracle checks")


EMULATED = os.environ.get("NHSTEP_EMU") == "1"


def _has_gpu():
    if EMULATED:
        return True
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


if EMULATED:
    # Development aid (tests/emu/README.md): run the `gpu` tests' logic against the kernels compiled for the host by tests/emu —
    # NHSTEP_EMU=1 python -m pytest tests/test_gpu_parity.py -m gpu.  Never set by the driver or by any default path.
    sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))
    import build_emu
    import oceananigans_jl_b200._lib as _L
    _L.LIB_PATH = build_emu.build(verbose=False)

    # the one-rank distributed driver: "device" buffers are host memory here, so torch sees them as CPU tensors
    import torch
    import oceananigans_jl_b200.distributed as _D

    import ctypes
    import numpy

    _D._DeviceMemory = lambda ptr, nbytes: numpy.ctypeslib.as_array((ctypes.c_ubyte * nbytes).from_address(ptr))
    _D._torch_device = lambda index: torch.device("cpu")
    torch.cuda.synchronize = lambda *a, **k: None


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)

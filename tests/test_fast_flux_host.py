"""Host-side guard for the tiled kernel's pointer-stepping WENO5 flux (csrc/tendency_generic.cuh, `tiled_adv_flux`, compile-time
switch TG_FASTFLUX): compiled as plain C++ it must equal the generic momentum / tracer flux functions bit for bit at every flux
point, on a triply periodic grid and with walls (tools/check_fast_flux_host.py).  No GPU involved; the CUDA build is untouched."""
import os
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("g++") is None, reason="needs g++")
def test_pointer_stepping_flux_equals_generic_flux_on_the_host():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "check_fast_flux_host.py")], capture_output=True, text=True,
                         timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    lines = [l for l in out.stdout.splitlines() if "flux points differ" in l]
    assert len(lines) == 24 and all(" 0 of " in l for l in lines), out.stdout

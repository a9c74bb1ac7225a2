"""The C ABI from a second language: tests/abi_smoke.c is compiled with gcc against include/nhstep.h and libnhstep.so, checks
the struct layouts that the ctypes binding and the Julia shim (julia/OceananigansB200Ext.jl) assume, and on a GPU steps a model."""
import ctypes as C
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def exe(tmp_path_factory):
    import __graft_entry__ as g
    libdir = os.path.join(ROOT, "oceananigans.jl_b200")
    if not os.path.exists(os.path.join(libdir, "libnhstep.so")):
        g.build()
    out = str(tmp_path_factory.mktemp("abi") / "abi_smoke")
    subprocess.run(["gcc", "-O1", "-Wall", "-Werror", "-std=c99", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "abi_smoke.c"), "-o", out, "-L", libdir, "-l:libnhstep.so", "-lm",
                    f"-Wl,-rpath,{libdir}"], check=True)
    return out


def _kv(text):
    return dict(line.split("=", 1) for line in text.splitlines() if "=" in line)


def test_struct_layouts_agree_between_c_ctypes_and_the_julia_shim(exe):
    import oceananigans_jl_b200 as oc
    L = oc._lib
    kv = _kv(subprocess.run([exe, "layout"], check=True, capture_output=True, text=True).stdout)
    assert int(kv["sizeof(nh_grid_desc)"]) == C.sizeof(L.GridDesc) == 96
    assert int(kv["offsetof(nh_grid_desc,x)"]) == L.GridDesc.x.offset == 40
    assert int(kv["offsetof(nh_grid_desc,z_faces)"]) == L.GridDesc.z_faces.offset == 88
    assert int(kv["sizeof(nh_model_desc)"]) == C.sizeof(L.ModelDesc)
    for f in ("ab2_chi", "n_tracers", "gravitational_acceleration", "coriolis_f", "nu", "kappa", "vertically_implicit",
              "smagorinsky_C", "amd_Cnu", "hydrostatic_pressure_anomaly", "exact_arithmetic"):
        assert int(kv[f"offsetof(nh_model_desc,{f})"]) == getattr(L.ModelDesc, f).offset, f
    # the numbers julia/OceananigansB200Ext.jl states for its NhModelDesc
    assert int(kv["offsetof(nh_model_desc,ab2_chi)"]) == 32
    assert int(kv["sizeof(nh_model_desc)"]) == 344
    assert int(kv["sizeof(nh_clock)"]) == C.sizeof(L.Clock) == 40
    # reconstruction_coefficients.jl:154-171 (Float32, Centered(4)), through C
    assert kv["centered4_f32"] == "-0.0833333358,0.583333313,0.583333313,-0.0833332539"


def test_c_program_fails_loudly_without_a_device(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    r = subprocess.run([exe, "step"], capture_output=True, text=True)
    assert r.returncode == 3 and "no CPU fallback" in r.stdout


@pytest.mark.gpu
def test_c_program_sets_and_steps_a_model(exe):
    r = subprocess.run([exe, "step"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "iteration=3" in r.stdout and "parent=22x22x22" in r.stdout

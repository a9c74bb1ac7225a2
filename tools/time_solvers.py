"""Standalone Poisson solvers on device-resident data (BASELINE.json configs[2], configs[3]):
   python tools/time_solvers.py [reps]
Prints ms per solve (CUDA events inside nh_poisson_time_solves) and the HBM rate for the 10 words per cell of SURVEY.md §8d."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oceananigans_jl_b200 as oc                      # noqa: E402
from oceananigans_jl_b200._lib import lib, check       # noqa: E402
from oceananigans_jl_b200.model import FFTBasedPoissonSolver, FourierTridiagonalPoissonSolver   # noqa: E402


def timed(solver, shape, FT, reps):
    tdt = torch.float64 if FT == np.float64 else torch.float32
    rhs = torch.randn(shape[::-1], dtype=tdt, device="cuda")
    rhs -= rhs.mean()
    phi = torch.empty_like(rhs)
    ms = C.c_float()
    check(lib().nh_poisson_time_solves(solver._h, C.c_void_p(rhs.data_ptr()), C.c_void_p(phi.data_ptr()), 2, C.byref(ms)))
    check(lib().nh_poisson_time_solves(solver._h, C.c_void_p(rhs.data_ptr()), C.c_void_p(phi.data_ptr()), reps, C.byref(ms)))
    return ms.value / reps


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    cases = []
    for FT in (np.float64, np.float32):
        for topo, name in (((oc.Periodic, oc.Periodic, oc.Bounded), "PPB"), ((oc.Periodic, oc.Periodic, oc.Periodic), "PPP")):
            for N in (256, 512):
                g = oc.RectilinearGrid(FT, size=(N, N, N), extent=(1, 1, 1), topology=topo)
                cases.append((f"FFTBasedPoissonSolver {N}^3 {name} {np.dtype(FT).name}", FFTBasedPoissonSolver(g), (N, N, N), FT))
        Nz = 256
        zf = -np.cos(np.linspace(0, np.pi / 2, Nz + 1))[::-1] if False else np.tanh(np.linspace(-2, 0, Nz + 1) * 1.0) / np.tanh(2.0)
        g = oc.RectilinearGrid(FT, size=(512, 512, Nz), x=(0, 1), y=(0, 1), z=np.ascontiguousarray(zf),
                               topology=(oc.Periodic, oc.Periodic, oc.Bounded))
        cases.append((f"FourierTridiagonalPoissonSolver 512x512x{Nz} stretched z {np.dtype(FT).name}",
                      FourierTridiagonalPoissonSolver(g), (512, 512, Nz), FT))
    for name, solver, shape, FT in cases:
        ms = timed(solver, shape, FT, reps)
        cells = shape[0] * shape[1] * shape[2]
        w = np.dtype(FT).itemsize
        print(f"{name:70s} {ms:8.3f} ms/solve  {cells / ms / 1e6:8.2f} Gcell/s   10w {10 * w * cells / ms / 1e9:7.2f} TB/s  "
              f"launches/solve {solver.launch_count() // (reps + 2 + 0)}")
        del solver


if __name__ == "__main__":
    main()

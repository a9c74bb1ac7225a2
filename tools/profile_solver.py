"""A few standalone FFTBasedPoissonSolver solves on device-resident data (development aid for ncu launch lists):
   python tools/profile_solver.py [N] [ppp|ppb] [f32]"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import oceananigans_jl_b200 as oc
from oceananigans_jl_b200._lib import lib, check
N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
topo = (oc.Periodic, oc.Periodic, oc.Bounded if "ppb" in sys.argv else oc.Periodic)
FT = np.float32 if "f32" in sys.argv else np.float64
g = oc.RectilinearGrid(FT, size=(N, N, N), extent=(1, 1, 1), topology=topo)
s = oc.FFTBasedPoissonSolver(g)
rhs = torch.randn((N, N, N), dtype=torch.float64 if FT == np.float64 else torch.float32, device="cuda")
phi = torch.empty_like(rhs)
ms = C.c_float()
check(lib().nh_poisson_time_solves(s._h, C.c_void_p(rhs.data_ptr()), C.c_void_p(phi.data_ptr()), 4, C.byref(ms)))
print(ms.value / 4, "ms per solve")

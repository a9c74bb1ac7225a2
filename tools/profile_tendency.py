"""Two steps of the 256^3 benchmark model in each arithmetic mode (exact first, then contracted): the target of the ncu
capture of tendency_fast_kernel<double, 0> and <double, 1>:
   ncu --set full --import-source on --clock-control none -k regex:tendency_fast --launch-skip 5 -c 2 \
       -o gpurun_out/tendency python tools/profile_tendency.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oceananigans_jl_b200 as oc
from bench import smooth_ics

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dt = 0.2 * (2 * np.pi / N) / 2.0
for exact in (True, False):
    g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(oc.Periodic,) * 3)
    m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic=exact)
    oc.set_model(m, **smooth_ics(N, 1234))
    oc.time_step(m, dt, n_steps=2)
    print("exact" if exact else "contracted", m.max_abs_divergence())

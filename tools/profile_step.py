"""Three steps of the 256^3 benchmark model (contracted arithmetic, the bench default): the target of the ncu capture of one
substep's seven kernels:
   ncu --set full --import-source on --clock-control none -k regex:"tendency_fast|pf_|pressure_correct_halo" \
       --launch-skip 28 -c 7 -o gpurun_out/step256 python tools/profile_step.py"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oceananigans_jl_b200 as oc
from bench import smooth_ics

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dt = 0.2 * (2 * np.pi / N) / 2.0
g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(oc.Periodic,) * 3)
m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic=False)
oc.set_model(m, **smooth_ics(N, 1234))
oc.time_step(m, dt, n_steps=3)
print(m.max_abs_divergence())

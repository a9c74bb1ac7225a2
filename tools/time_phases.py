"""Quick phase timing of one model configuration (development aid): python tools/time_phases.py [N] [steps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oceananigans_jl_b200 as oc
from bench import smooth_ics

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
FT = np.float32 if (len(sys.argv) > 3 and sys.argv[3] == "f32") else np.float64
g = oc.RectilinearGrid(FT, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(oc.Periodic,) * 3)
m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5))
oc.set_model(m, **smooth_ics(N, 1234))
dt = 0.2 * (2 * np.pi / N) / 2.0
oc.time_step(m, dt, n_steps=3)
acc = {}
for _ in range(steps):
    for k, v in oc.time_step_profiled(m, dt).items():
        acc[k] = acc.get(k, 0.0) + v / steps
print({k: round(v / 3, 4) for k, v in acc.items()}, "ms per substep; step", round(sum(acc.values()), 3), "ms; div", m.max_abs_divergence())

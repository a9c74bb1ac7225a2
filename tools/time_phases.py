"""Quick phase timing of one model configuration (development aid):
   python tools/time_phases.py [N] [steps] [f32] [ppb | les [vitd] | smag | amd | c4]
   ppb: (Periodic, Periodic, Bounded) WENO5, no tracers;  les: ppb + buoyancy tracer b + ScalarDiffusivity + FPlane;
   smag / amd: les with SmagorinskyLilly / AnisotropicMinimumDissipation and the ocean-LES wall conditions (wind-stress flux
   on u, no-slip bottom, surface buoyancy flux and bottom buoyancy gradient) — the widened rows of DESIGN.md §7b"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oceananigans_jl_b200 as oc
from bench import smooth_ics

N = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
FT = np.float32 if "f32" in sys.argv else np.float64
eddy = "smag" in sys.argv or "amd" in sys.argv
bounded = "ppb" in sys.argv or "les" in sys.argv or eddy
topo = (oc.Periodic, oc.Periodic, oc.Bounded if bounded else oc.Periodic)
if "c4" in sys.argv:
    # BASELINE.json configs[3]: N × N × N/2, tanh-stretched Bounded z (test_time_stepping.jl:439-445), Fourier-tridiagonal solve
    Nz = N // 2
    kk = np.arange(1, Nz + 2)
    zf = np.tanh(1.3 * (2 * (kk - 1) / Nz - 1)) / np.tanh(1.3)
    topo = (oc.Periodic, oc.Periodic, oc.Bounded)
    g = oc.RectilinearGrid(FT, size=(N, N, Nz), x=(0, 2 * np.pi), y=(0, 2 * np.pi), z=zf, topology=topo)
    bounded = True
else:
    Nz = N
    g = oc.RectilinearGrid(FT, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=topo)
kw = {}
if "les" in sys.argv:
    kw = dict(buoyancy=oc.BuoyancyTracer(), tracers=("b",), closure=oc.ScalarDiffusivity(1e-4, 1e-4, vertically_implicit="vitd" in sys.argv), coriolis=oc.FPlane(1e-4))
if eddy:
    bcs = {"u": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(-1e-4), bottom=oc.ValueBoundaryCondition(0.0)),
           "v": oc.FieldBoundaryConditions(bottom=oc.ValueBoundaryCondition(0.0)),
           "b": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(1e-8), bottom=oc.GradientBoundaryCondition(1e-5))}
    kw = dict(buoyancy=oc.BuoyancyTracer(), tracers=("b",), coriolis=oc.FPlane(1e-4), boundary_conditions=bcs,
              closure=oc.SmagorinskyLilly() if "smag" in sys.argv else oc.AnisotropicMinimumDissipation())
m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic="fma" not in sys.argv, **kw)
ics = smooth_ics(N, 1234, size=(N, N, Nz))
if bounded:
    ics["w"] = np.concatenate([ics["w"], ics["w"][:, :, :1]], axis=2)
    ics["w"][:, :, 0] = 0
    ics["w"][:, :, -1] = 0
oc.set_model(m, **ics)
dt = 0.2 * (2 * np.pi / N) / 2.0
oc.time_step(m, dt, n_steps=3)
acc = {}
for _ in range(steps):
    for k, v in oc.time_step_profiled(m, dt).items():
        acc[k] = acc.get(k, 0.0) + v / steps
print({k: round(v / 3, 4) for k, v in acc.items()}, "ms per substep; step", round(sum(acc.values()), 3), "ms; div", m.max_abs_divergence())

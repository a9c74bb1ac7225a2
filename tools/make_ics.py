"""Initial conditions for tools/dump_reference.jl (the third-party value-parity recipe of SURVEY.md §8c):

    python tools/make_ics.py [N]               writes ics_{u,v,w}_N.bin: the bench / test initial state (bench.smooth_ics, seed 1234)
                                               as raw little-endian Float64, column-major (x fastest) like a Julia Array
    N=32 julia tools/dump_reference.jl         (any machine with Julia + Oceananigans v0.96.19) writes ref_{u,v,w,p}_N.bin
    python tools/make_ics.py [N] --compare     steps the oracle (and, on a GPU box, libnhstep.so) 100 times from the same files
                                               and prints the rel. L2 differences against ref_*.bin

Neither Julia nor the reference's remote regression files are available in the build image, so the ref_*.bin of the last step
are not committed; everything on this side of the recipe runs here.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import smooth_ics   # noqa: E402


def write_ics(N):
    for name, a in smooth_ics(N, 1234).items():
        np.asfortranarray(a, dtype="<f8").ravel(order="F").tofile(f"ics_{name}_{N}.bin")
        print(f"wrote ics_{name}_{N}.bin")


def read(name, N):
    return np.fromfile(name, dtype="<f8").reshape((N, N, N), order="F")


def compare(N):
    import oracle as o
    ics = {n: read(f"ics_{n}_{N}.bin", N) for n in "uvw"}
    dt = 0.2 * (2 * np.pi / N) / 2.0
    topo = ("Periodic",) * 3
    og = o.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=topo)
    om = o.NonhydrostaticModel(og, advection=o.WENO(np.float64, 5))
    o.set_model(om, **ics)
    for _ in range(100):
        o.time_step(om, dt)
    mine = {"oracle": {"u": om.velocities[0].interior, "v": om.velocities[1].interior, "w": om.velocities[2].interior,
                       "p": om.pNHS.interior}}
    try:
        import oceananigans_jl_b200 as oc
        if oc.device_count() >= 1:
            g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=topo)
            m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5))
            oc.set_model(m, **ics)
            oc.time_step(m, dt, n_steps=100)
            mine["libnhstep"] = {n: m.velocities[n].interior() for n in "uvw"} | {"p": m.pressures.pNHS.interior()}
    except Exception as e:                           # no device here: the oracle column alone
        print("libnhstep.so not run:", e)
    for who, fields in mine.items():
        for n, a in fields.items():
            path = f"ref_{n}_{N}.bin"
            if not os.path.exists(path):
                print(f"{who} {n}: {path} missing (run tools/dump_reference.jl with Julia first)")
                continue
            r = read(path, N)
            print(f"{who:10s} {n}: rel. L2 vs the reference = {np.linalg.norm((a - r).ravel()) / np.linalg.norm(r.ravel()):.3e}")


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    N = int(args[0]) if args else 32
    compare(N) if "--compare" in sys.argv else write_ics(N)

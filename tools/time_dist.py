"""Development timing of the y-slab distributed step:  torchrun --nproc-per-node R tools/time_dist.py Nx Ny Nz [steps] [f32]
(also runs with R = 1 without torchrun).  Prints ms/step (max over ranks) and a per-piece breakdown from CUDA events."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C                                     # noqa: E402
import oceananigans_jl_b200 as oc                      # noqa: E402
from oceananigans_jl_b200 import distributed as D      # noqa: E402
from oceananigans_jl_b200 import _lib as L             # noqa: E402
from oceananigans_jl_b200._lib import lib, check       # noqa: E402


def main():
    Nx, Ny, Nz = (int(a) for a in sys.argv[1:4])
    steps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
    FT = np.float32 if "f32" in sys.argv else np.float64
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    arch = D.Distributed(D.Partition(y=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=(Nx, Ny, Nz), x=(0, 2 * np.pi), y=(0, 2 * np.pi), z=(0, 2 * np.pi),
                                         advection=oc.WENO(FT, 5), device=local, p2p="p2p" in sys.argv)
    rng = np.random.default_rng(1)
    x, y, z = np.arange(Nx) / Nx, (m.j0 + np.arange(m.NyL)) / Ny, np.arange(Nz) / Nz     # this rank's rows only
    ics = {}
    for name in "uvw":
        k, ph = rng.integers(1, 4, 3), rng.uniform(0, 6.28, 3)
        ics[name] = (np.cos(6.28 * k[0] * x + ph[0])[:, None, None] * np.cos(6.28 * k[1] * y + ph[1])[None, :, None]
                     * np.cos(6.28 * k[2] * z + ph[2])[None, None, :]).astype(FT)
    m.set_local(**ics)
    dt = 0.2 * (2 * np.pi / Nx) / 2.0

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    m.time_step(dt, n_steps=3)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    m.time_step(dt, n_steps=steps)
    e1.record()
    sync()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)

    # pieces of one stage, each timed alone (events around it, synchronised on both sides)
    h, b = m._h, m._buf
    pieces = {}

    def timed(name, fn, reps=5):
        sync()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        c.record()
        sync()
        pieces[name] = a.elapsed_time(c) / reps

    sdt = C.c_double()
    timed("tendency_substep", lambda: check(lib().nh_dist_tendency_substep(h, dt, 1, 0, C.byref(sdt))))
    timed("barrier", m._barrier if m.p2p else (lambda: None))
    timed("exchange_v_1row", lambda: m.exchange(1 << L.FIELD_V, 1))
    timed("poisson_forward(x,z)", lambda: check(lib().nh_dist_poisson(h, 0, dt)))
    timed("all_to_all", lambda: D.transpose_blocks(b[L.DIST_SPECTRAL_B], b[L.DIST_SPECTRAL_A], world, arch.group))
    timed("poisson_solve_y", lambda: check(lib().nh_dist_poisson(h, 1, dt)))
    timed("poisson_backward(z,x)", lambda: check(lib().nh_dist_poisson(h, 2, dt)))
    timed("exchange_p_1row", lambda: m.exchange(1 << L.FIELD_PNHS, 1))
    timed("pressure_correct", lambda: check(lib().nh_dist_pressure_correct(h, dt)))
    timed("exchange_uvw_3rows", lambda: m.exchange(m._uvw, 3))
    if rank == 0:
        cells = Nx * Ny * Nz
        print(f"p2p={m.p2p} R={world} global {Nx}x{Ny}x{Nz} {np.dtype(FT).name}: {ms.item():.3f} ms/step, "
              f"{cells / (ms.item() * 1e-3):.3e} cell-updates/s; per stage pieces (ms):")
        for k, v in pieces.items():
            print(f"   {k:24s} {v:8.3f}")
        print(f"   sum x3 = {3 * (sum(pieces.values()) + pieces['all_to_all']):.3f} ms")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

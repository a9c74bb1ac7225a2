"""Timing of a complete-path model on y-slabs (weak scaling: N x N*R x Nz cells over R GPUs):
   torchrun --nproc-per-node R tools/time_dist_generic.py [N] [steps] [les | smag | stretched]
(also runs with R = 1 without torchrun).  Prints ms/step (max over ranks) and cell-updates/s of the whole job."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oceananigans_jl_b200 as oc                      # noqa: E402
from oceananigans_jl_b200 import distributed as D      # noqa: E402


def main():
    nums = [int(a) for a in sys.argv[1:] if a.isdigit()]
    N = nums[0] if nums else 256
    steps = nums[1] if len(nums) > 1 else 5
    FT = np.float32 if "f32" in sys.argv else np.float64
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    Nz = N // 2 if "stretched" in sys.argv else N
    if "stretched" in sys.argv:
        k = np.arange(Nz + 1)
        z = np.tanh(1.3 * (2 * k / Nz - 1)) / np.tanh(1.3)
    else:
        z = (0, 2 * np.pi)
    bcs = {"u": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(-1e-4), bottom=oc.ValueBoundaryCondition(0.0)),
           "b": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(1e-8), bottom=oc.GradientBoundaryCondition(1e-5))}
    kw = dict(buoyancy=oc.BuoyancyTracer(), tracers=("b",), coriolis=oc.FPlane(1e-4), boundary_conditions=bcs,
              closure=oc.SmagorinskyLilly() if "smag" in sys.argv else oc.ScalarDiffusivity(1e-4, 1e-4))
    arch = D.Distributed(D.Partition(y=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=(N, N * world, Nz), x=(0, 2 * np.pi), y=(0, 2 * np.pi * world), z=z,
                                         advection=oc.WENO(FT, 5), device=local,
                                         topology=(oc.Periodic, oc.Periodic, oc.Bounded), **kw)
    rng = np.random.default_rng(1)
    x, y, zc = np.arange(N) / N, (m.j0 + np.arange(m.NyL)) / N, np.arange(Nz + 1) / Nz
    ics = {}
    for name in ("u", "v", "w", "b"):
        kk, ph = rng.integers(1, 4, 3), rng.uniform(0, 6.28, 3)
        nz = Nz + 1 if name == "w" else Nz
        f = (np.cos(6.28 * kk[0] * x + ph[0])[:, None, None] * np.cos(6.28 * kk[1] * y + ph[1])[None, :, None]
             * np.sin(3.14159 * kk[2] * zc[:nz])[None, None, :]).astype(FT)
        ics[name] = f
    m.set_local(**ics)
    dt = 0.05 * (2 * np.pi / N) / 2.0

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    m.time_step(dt, n_steps=2)
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    m.time_step(dt, n_steps=steps)
    e1.record()
    sync()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    div = m.max_abs_divergence()
    if rank == 0:
        cells = N * N * world * Nz
        print({"ranks": world, "cells_per_gpu": N * N * Nz, "ms_per_step": round(float(ms.item()), 3),
               "cell_updates_per_s": cells / (float(ms.item()) * 1e-3), "max_div": div, "config": [a for a in sys.argv[1:] if not a.isdigit()]})
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

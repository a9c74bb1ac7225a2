"""GPU: Partition(x=R) — x-slabs, the reference's default slab direction (distributed_architectures.jl:12-118) — against one GPU.

The kernels partition in y; an x-slab run is the y-slab run of the mirrored problem (oceananigans.jl_b200/distributed.py).  Written
after the round's GPU budget was spent: verified with tests/emu (one process per rank on host cores), first run on GPUs pending; kept
in a late-collected file of its own for that reason.  R = 2, 4 need that many GPUs.
"""
import socket

import numpy as np
import pytest

from test_gpu_distributed import (_check, _generic_check, _generic_ics, _generic_kwargs, _generic_single, _generic_size, _ics,
                                  _init_rank, _ranks_available, _single_gpu, _zfaces, GENERIC)

pytestmark = pytest.mark.gpu


def _swap(a):
    return np.ascontiguousarray(np.asarray(a).transpose(1, 0, 2))


def _x_slab_fused(rank, world, size, FT, timestepper, steps, dt, out, p2p):
    """The fused triply periodic model of _single_gpu (x ∈ (0, 1), y ∈ (0, 2)) partitioned in x."""
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    arch = D.Distributed(D.Partition(x=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=size, x=(0, 1), y=(0, 2), z=(0, 1), halo=(3, 3, 3),
                                         advection=oc.WENO(FT, 5), timestepper=timestepper, device=rank, p2p=p2p)
    assert m.mirrored and m.local_interior("u").shape == (size[0] // world, size[1], size[2])
    m.set(**_ics(size, FT))
    div0 = m.max_abs_divergence()
    m.time_step(dt, n_steps=steps)
    res = {n: m.gather(n) for n in ("u", "v", "w", "pNHS")}
    res["div0"], res["div"] = div0, m.max_abs_divergence()
    res["time"], res["iteration"] = m.clock.time, m.clock.iteration
    if rank == 0:
        out.update(res)


def _x_slab_generic(rank, world, name, FT, timestepper, steps, dt, out):
    """A complete-path model (tracers, buoyancy, Coriolis, closure, walls — test_gpu_distributed.GENERIC) partitioned in x."""
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    size, bounded, stretched = _x_size(name, world)
    topo = (oc.Periodic, oc.Periodic, oc.Bounded if bounded else oc.Periodic)
    kw = _generic_kwargs(oc, name, FT)
    arch = D.Distributed(D.Partition(x=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=size, x=(0, 1), y=(0, 2), z=_zfaces(size[2], stretched), halo=(3, 3, 3),
                                         advection=oc.WENO(FT, 5), timestepper=timestepper, device=rank, topology=topo, **kw)
    assert m.generic and m.mirrored
    m.set(**_generic_ics(size, FT, bounded, kw["tracers"]))
    m.time_step(dt, n_steps=steps)
    res = {n: m.gather(n) for n in ("u", "v", "w", "pNHS") + tuple(kw["tracers"])}
    res["time"], res["iteration"], res["div"] = m.clock.time, m.clock.iteration, m.max_abs_divergence()
    if rank == 0:
        out.update(res)


def _x_size(name, world):
    size, bounded, stretched = GENERIC[name]
    return (max(size[0], 8 * world), size[1], size[2]), bounded, stretched


def _x_generic_reference(name, FT, timestepper, steps, dt, world):
    import oceananigans_jl_b200 as oc
    size, bounded, stretched = _x_size(name, world)
    topo = (oc.Periodic, oc.Periodic, oc.Bounded if bounded else oc.Periodic)
    g = oc.RectilinearGrid(FT, size=size, x=(0, 1), y=(0, 2), z=_zfaces(size[2], stretched), halo=(3, 3, 3), topology=topo)
    kw = _generic_kwargs(oc, name, FT)
    m = oc.NonhydrostaticModel(g, advection=oc.WENO(FT, 5), timestepper=timestepper, **kw)
    oc.set_model(m, **_generic_ics(size, FT, bounded, kw["tracers"]))
    oc.time_step(m, dt, n_steps=steps)
    fields = {n: m.velocities[n].interior() for n in "uvw"} | {n: m.tracers[n].interior() for n in kw["tracers"]}
    fields["pNHS"] = m.pressures.pNHS.interior()
    return fields | {"time": m.clock.time, "iteration": m.clock.iteration}


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_one_rank_x_slab_matches_single_gpu(timestepper, p2p):
    size, FT, steps, dt = (32, 16, 16), np.float64, 10, 2e-3
    ref = _single_gpu(size, FT, timestepper, steps, dt)
    out = {}
    _x_slab_fused(0, 1, size, FT, timestepper, steps, dt, out, p2p)
    _check(out, ref, FT, steps, dt)


def _fused_worker(rank, world, port, size, FT, timestepper, steps, dt, path, p2p):
    import os
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    _init_rank(rank, world)
    try:
        out = {}
        _x_slab_fused(rank, world, size, FT, timestepper, steps, dt, out, p2p)
        if rank == 0:
            np.savez(path, **out)
    finally:
        dist.destroy_process_group()


def _generic_worker(rank, world, port, name, FT, timestepper, steps, dt, path):
    import os
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    _init_rank(rank, world)
    try:
        out = {}
        _x_slab_generic(rank, world, name, FT, timestepper, steps, dt, out)
        if rank == 0:
            np.savez(path, **out)
    finally:
        dist.destroy_process_group()


def _port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("FT,timestepper", [(np.float64, "RungeKutta3"), (np.float32, "QuasiAdamsBashforth2")])
def test_x_slabs_match_single_gpu(world, FT, timestepper, p2p, tmp_path):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    steps, dt = 10, 2e-3
    size = (max(32, 8 * world), 16, 32)
    ref = _single_gpu(size, FT, timestepper, steps, dt)
    path = str(tmp_path / "x.npz")
    mp.spawn(_fused_worker, args=(world, _port(), size, FT, timestepper, steps, dt, path, p2p), nprocs=world, join=True)
    res = dict(np.load(path))
    _check({k: (v if v.ndim else v.item()) for k, v in res.items()}, ref, FT, steps, dt)


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("name", list(GENERIC))
def test_complete_path_x_slabs_match_single_gpu(name, world, tmp_path):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    FT, timestepper, steps, dt = np.float64, "RungeKutta3", 10, 2e-3
    ref = _x_generic_reference(name, FT, timestepper, steps, dt, world)
    path = str(tmp_path / "xg.npz")
    mp.spawn(_generic_worker, args=(world, _port(), name, FT, timestepper, steps, dt, path), nprocs=world, join=True)
    res = dict(np.load(path))
    _generic_check({k: (v if v.ndim else v.item()) for k, v in res.items()}, ref, FT, name)

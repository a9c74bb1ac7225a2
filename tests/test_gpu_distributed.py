"""Multi-GPU (y-slab) time step against the single-GPU model and the oracle.

R = 1 exercises every distributed kernel (blocked spectral layout, z/y passes, halo pack/unpack, ywrap = 0 paths) on one
GPU with the messages degenerating to copies; R = 2, 4 need that many GPUs (``gpurun --gpus N``) and run one process per
GPU over NCCL.  Tolerance: F64 relative L2 ≤ 1e-12 (the FFT passes run in a different order than on one GPU, so the
results are not bit-identical), F32 ≤ 1e-5.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

# tests/emu (NHSTEP_EMU=1, development aid): one process per rank on host cores, the peers' buffers as files under /dev/shm
EMULATED = os.environ.get("NHSTEP_EMU") == "1"


def _ranks_available():
    return 8 if EMULATED else torch.cuda.device_count()


def _init_rank(rank, world):
    import torch.distributed as dist
    if EMULATED:
        os.environ["NHSTEP_EMU_SHM"] = "1"          # before the first allocation of this process
        import conftest  # noqa: F401                 (switches this process to the host stand-in as it did for the parent)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        return
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))


def _ics(size, FT, seed=3):
    rng = np.random.default_rng(seed)
    Nx, Ny, Nz = size
    x = (np.arange(Nx) + 0.5) / Nx
    y = (np.arange(Ny) + 0.5) / Ny
    z = (np.arange(Nz) + 0.5) / Nz
    X, Y, Z = np.meshgrid(x, y, z, indexing="ij")
    u = np.sin(2 * np.pi * X) * np.cos(2 * np.pi * Y) * np.cos(4 * np.pi * Z) + 0.1 * rng.standard_normal(size)
    v = -np.cos(2 * np.pi * X) * np.sin(2 * np.pi * Y) * np.cos(2 * np.pi * Z) + 0.1 * rng.standard_normal(size)
    w = 0.3 * np.sin(4 * np.pi * X) * np.sin(2 * np.pi * Y) + 0.1 * rng.standard_normal(size)
    return {k: a.astype(FT) for k, a in (("u", u), ("v", v), ("w", w))}


def _single_gpu(size, FT, timestepper, steps, dt):
    import oceananigans_jl_b200 as oc
    g = oc.RectilinearGrid(FT, size=size, x=(0, 1), y=(0, 2), z=(0, 1), halo=(3, 3, 3),
                           topology=(oc.Periodic, oc.Periodic, oc.Periodic))
    m = oc.NonhydrostaticModel(g, advection=oc.WENO(FT, 5), timestepper=timestepper)
    oc.set_model(m, **_ics(size, FT))
    oc.time_step(m, dt, n_steps=steps)
    return {n: m.velocities[n].interior() for n in "uvw"} | {"pNHS": m.pressures.pNHS.interior(), "time": m.clock.time}


def _distributed(rank, world, size, FT, timestepper, steps, dt, out, p2p=False, exact=True):
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    arch = D.Distributed(D.Partition(y=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=size, x=(0, 1), y=(0, 2), z=(0, 1), halo=(3, 3, 3),
                                         advection=oc.WENO(FT, 5), timestepper=timestepper, device=rank, p2p=p2p,
                                         exact_arithmetic=exact)
    m.set(**_ics(size, FT))
    div0 = m.max_abs_divergence()
    m.time_step(dt, n_steps=steps)
    res = {n: m.gather(n) for n in ("u", "v", "w", "pNHS")}
    res["div0"], res["div"] = div0, m.max_abs_divergence()
    res["time"], res["iteration"] = m.clock.time, m.clock.iteration
    if rank == 0:
        out.update(res)
    return res


def _rel(a, b):
    return float(np.linalg.norm((a.astype(np.float64) - b.astype(np.float64)).ravel()) / max(np.linalg.norm(b.astype(np.float64).ravel()), 1e-300))


def _check(res, ref, FT, steps, dt):
    tol = 1e-12 if FT == np.float64 else 1e-5
    for n in ("u", "v", "w", "pNHS"):
        assert _rel(res[n], ref[n]) <= tol, (n, _rel(res[n], ref[n]))
    scale = 1e-13 if FT == np.float64 else 1e-4
    assert res["div0"] < scale * 100 and res["div"] < scale * 100
    assert res["iteration"] == steps and res["time"] == ref["time"]


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("size", [(32, 32, 16), (16, 64, 32)])
def test_one_rank_distributed_path_matches_single_gpu(FT, timestepper, size, p2p):
    steps, dt = 10, 2e-3
    ref = _single_gpu(size, FT, timestepper, steps, dt)
    out = {}
    _distributed(0, 1, size, FT, timestepper, steps, dt, out, p2p)
    _check(out, ref, FT, steps, dt)


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "contracted"])
def test_one_rank_distributed_path_matches_oracle(exact):
    import oracle as o
    size, steps, dt, FT = (16, 32, 16), 10, 2e-3, np.float64
    grid = o.RectilinearGrid(FT, size=size, x=(0, 1), y=(0, 2), z=(0, 1), halo=(3, 3, 3),
                             topology=(o.Periodic, o.Periodic, o.Periodic))
    m = o.NonhydrostaticModel(grid, advection=o.WENO(FT, 5), timestepper="RungeKutta3")
    o.set_model(m, **_ics(size, FT))
    for _ in range(steps):
        o.time_step(m, dt)
    ref = {n: m.velocities[a].interior for a, n in enumerate("uvw")} | {"pNHS": m.pNHS.interior}
    out = {}
    _distributed(0, 1, size, FT, "RungeKutta3", steps, dt, out, exact=exact)
    for n in ("u", "v", "w", "pNHS"):
        assert _rel(out[n], ref[n]) <= 1e-12, (n, _rel(out[n], ref[n]))


def _worker(rank, world, port, size, FT, timestepper, steps, dt, path, p2p):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    _init_rank(rank, world)
    try:
        out = {}
        _distributed(rank, world, size, FT, timestepper, steps, dt, out, p2p)
        if rank == 0:
            np.savez(path, **out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("FT,timestepper,size", [(np.float64, "RungeKutta3", (32, 64, 16)),
                                                 (np.float64, "QuasiAdamsBashforth2", (64, 32, 32)),
                                                 (np.float32, "RungeKutta3", (32, 64, 32))])
def test_multi_gpu_matches_single_gpu(world, FT, timestepper, size, p2p, tmp_path):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    steps, dt = 10, 2e-3
    size = (size[0], max(size[1], 8 * world), size[2])          # the fused sweep needs slabs of at least 2·H = 6 rows
    ref = _single_gpu(size, FT, timestepper, steps, dt)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    path = str(tmp_path / "dist.npz")
    mp.spawn(_worker, args=(world, port, size, FT, timestepper, steps, dt, path, p2p), nprocs=world, join=True)
    res = dict(np.load(path))
    _check(res, ref, FT, steps, dt)


# ---- complete-path models on a slab (tracers, buoyancy, closures, boundary conditions, Bounded / stretched z) -----------------
# DistributedFFTBasedPoissonSolver / DistributedFourierTridiagonalPoissonSolver rows of SURVEY.md §8 a32: the y halo fill pulls
# the neighbours' rows over NVLink, the solve transposes through peer stores, the host loop is the single-GPU one.

def _zfaces(Nz, stretched):
    if not stretched:
        return (0, 1)
    k = np.arange(Nz + 1)
    return 0.5 * (1 + np.tanh(1.3 * (2 * k / Nz - 1)) / np.tanh(1.3))


GENERIC = {
    # name: (size, z topology Bounded?, stretched, model kwargs builder)
    "ppp-tracer": ((32, 32, 16), False, False),
    "ppb-les": ((32, 32, 16), True, False),
    "ppb-stretched-smag": ((16, 32, 12), True, True),
}


def _generic_kwargs(oc, name, FT):
    if name == "ppp-tracer":
        return dict(tracers=("c",), closure=oc.ScalarDiffusivity(1e-3, 2e-3), coriolis=oc.FPlane(0.5))
    bcs = {"u": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(-1e-2), bottom=oc.ValueBoundaryCondition(0.0)),
           "b": oc.FieldBoundaryConditions(top=oc.FluxBoundaryCondition(1e-3), bottom=oc.GradientBoundaryCondition(0.1))}
    if name == "ppb-les":
        return dict(tracers=("b",), buoyancy=oc.BuoyancyTracer(), coriolis=oc.FPlane(0.3), boundary_conditions=bcs,
                    closure=oc.ScalarDiffusivity(1e-3, 1e-3, vertically_implicit=True))
    return dict(tracers=("b",), buoyancy=oc.BuoyancyTracer(), boundary_conditions=bcs, closure=oc.SmagorinskyLilly())


def _generic_ics(size, FT, bounded, names):
    ics = _ics(size, FT)
    if bounded:
        w = np.concatenate([ics["w"], ics["w"][:, :, :1]], axis=2)
        w[:, :, 0] = 0
        w[:, :, -1] = 0
        ics["w"] = w
    rng = np.random.default_rng(11)
    for n in names:
        ics[n] = (0.2 * rng.standard_normal(size)).astype(FT)
    return ics


def _generic_size(name, world):
    """The configuration's size with Ny raised so that every slab keeps at least 8 rows (a slab needs 2·H = 6)."""
    size, bounded, stretched = GENERIC[name]
    return (size[0], max(size[1], 8 * world), size[2]), bounded, stretched


def _generic_single(name, FT, timestepper, steps, dt, world=1):
    import oceananigans_jl_b200 as oc
    size, bounded, stretched = _generic_size(name, world)
    topo = (oc.Periodic, oc.Periodic, oc.Bounded if bounded else oc.Periodic)
    g = oc.RectilinearGrid(FT, size=size, x=(0, 1), y=(0, 2), z=_zfaces(size[2], stretched), halo=(3, 3, 3), topology=topo)
    kw = _generic_kwargs(oc, name, FT)
    m = oc.NonhydrostaticModel(g, advection=oc.WENO(FT, 5), timestepper=timestepper, **kw)
    oc.set_model(m, **_generic_ics(size, FT, bounded, kw["tracers"]))
    oc.time_step(m, dt, n_steps=steps)
    fields = {n: m.velocities[n].interior() for n in "uvw"} | {n: m.tracers[n].interior() for n in kw["tracers"]}
    fields["pNHS"] = m.pressures.pNHS.interior()
    return fields | {"time": m.clock.time, "iteration": m.clock.iteration}


def _generic_distributed(rank, world, name, FT, timestepper, steps, dt, out):
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    size, bounded, stretched = _generic_size(name, world)
    topo = (oc.Periodic, oc.Periodic, oc.Bounded if bounded else oc.Periodic)
    kw = _generic_kwargs(oc, name, FT)
    arch = D.Distributed(D.Partition(y=world))
    m = D.DistributedNonhydrostaticModel(arch, FT, size=size, x=(0, 1), y=(0, 2), z=_zfaces(size[2], stretched), halo=(3, 3, 3),
                                         advection=oc.WENO(FT, 5), timestepper=timestepper, device=rank, topology=topo, **kw)
    assert m.generic
    m.set(**_generic_ics(size, FT, bounded, kw["tracers"]))
    m.time_step(dt, n_steps=steps)
    res = {n: m.gather(n) for n in ("u", "v", "w", "pNHS") + tuple(kw["tracers"])}
    res["time"], res["iteration"], res["div"] = m.clock.time, m.clock.iteration, m.max_abs_divergence()
    if rank == 0:
        out.update(res)


def _generic_check(res, ref, FT, name):
    # stretched z: the same Fourier-tridiagonal algorithm on both sides; regular Bounded z: the slab run solves with the
    # tridiagonal sweep where one GPU uses the cosine transform, triply periodic: FFT passes in a different order
    tol = 1e-12 if FT == np.float64 else 1e-5
    for n in ref:
        if n in ("time", "iteration"):
            continue
        assert _rel(res[n], ref[n]) <= tol, (name, n, _rel(res[n], ref[n]))
    assert res["iteration"] == ref["iteration"] and res["time"] == ref["time"]
    assert res["div"] < (1e-11 if FT == np.float64 else 1e-3)


@pytest.mark.parametrize("FT", [np.float64, np.float32])
@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
@pytest.mark.parametrize("name", list(GENERIC))
def test_one_rank_complete_path_slab_matches_single_gpu(name, timestepper, FT):
    steps, dt = 10, 2e-3
    ref = _generic_single(name, FT, timestepper, steps, dt)
    out = {}
    _generic_distributed(0, 1, name, FT, timestepper, steps, dt, out)
    _generic_check(out, ref, FT, name)


def _generic_worker(rank, world, port, name, FT, timestepper, steps, dt, path):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    _init_rank(rank, world)
    try:
        out = {}
        _generic_distributed(rank, world, name, FT, timestepper, steps, dt, out)
        if rank == 0:
            np.savez(path, **out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("FT,timestepper", [(np.float64, "RungeKutta3"), (np.float32, "QuasiAdamsBashforth2")])
@pytest.mark.parametrize("name", list(GENERIC))
def test_multi_gpu_complete_path_slab_matches_single_gpu(name, FT, timestepper, world, tmp_path):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    steps, dt = 10, 2e-3
    ref = _generic_single(name, FT, timestepper, steps, dt, world)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    path = str(tmp_path / "dist.npz")
    mp.spawn(_generic_worker, args=(world, port, name, FT, timestepper, steps, dt, path), nprocs=world, join=True)
    res = dict(np.load(path))
    _generic_check(res, ref, FT, name)


# ---- bit-identical transpose round trip (test/test_distributed_transpose.jl:13-54) ---------------------------------------------
def _transpose_round_trip(rank, world, size, p2p):
    """Fill the z-local spectral buffer A with rank- and position-dependent values, transpose z → y → z through the library's
    peer-store kernels (or the NCCL all-to-all), and require A back bit for bit; B in between must hold, for every source
    rank s, block s = the block `rank` of rank s's A."""
    import oceananigans_jl_b200 as oc
    from oceananigans_jl_b200 import distributed as D
    from oceananigans_jl_b200 import _lib as L
    arch = D.Distributed(D.Partition(y=world))
    m = D.DistributedNonhydrostaticModel(arch, np.float64, size=size, x=(0, 1), y=(0, 1), z=(0, 1), advection=oc.WENO(order=5),
                                         device=rank, p2p=p2p)
    A, B = m._buf[L.DIST_SPECTRAL_A], m._buf[L.DIST_SPECTRAL_B]
    n = A.numel()
    chunk = n // world
    gen = torch.Generator(device="cpu").manual_seed(100 + rank)
    a0 = torch.view_as_complex(torch.randn(n, 2, dtype=torch.float64, generator=gen)).to(A.device)
    A.copy_(a0)
    B.zero_()
    torch.cuda.synchronize()
    m.transpose(0)
    torch.cuda.synchronize()
    # what B must hold: block s = block `rank` of rank s's A (regenerated from rank s's seed)
    for s_ in range(world):
        gs = torch.Generator(device="cpu").manual_seed(100 + s_)
        as_ = torch.view_as_complex(torch.randn(n, 2, dtype=torch.float64, generator=gs))
        want = as_[rank * chunk:(rank + 1) * chunk].to(A.device)
        assert torch.equal(B[s_ * chunk:(s_ + 1) * chunk], want), (rank, s_)
    A.zero_()
    torch.cuda.synchronize()
    m.transpose(1)
    torch.cuda.synchronize()
    assert torch.equal(A, a0), "A -> B -> A must be the identity, bit for bit"
    return True


def _transpose_worker(rank, world, port, size, p2p):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    _init_rank(rank, world)
    try:
        _transpose_round_trip(rank, world, size, p2p)
    finally:
        dist.destroy_process_group()


def test_transpose_round_trip_is_bit_identical_on_one_rank():
    assert _transpose_round_trip(0, 1, (32, 32, 16), True)


@pytest.mark.parametrize("p2p", [False, True], ids=["nccl", "p2p"])
@pytest.mark.parametrize("world", [2, 4, 8])
def test_transpose_round_trip_is_bit_identical(world, p2p):
    if _ranks_available() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_transpose_worker, args=(world, port, (32, 16 * world, 16), p2p), nprocs=world, join=True)

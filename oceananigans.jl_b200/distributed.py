"""Multi-GPU NonhydrostaticModel: one process per GPU, slab partition (y-slabs; x-slabs as their mirror image), NCCL over NVLink
through ``torch.distributed``.

Mirrors the reference's distributed architecture for the hot path:
  Distributed(GPU(); partition = Partition(1, R, 1))    src/DistributedComputations/distributed_architectures.jl:12-118,182-285
  fill_halo_regions! of distributed fields              src/DistributedComputations/halo_communication.jl:114-366
  DistributedFFTBasedPoissonSolver + transposes         src/DistributedComputations/distributed_fft_based_poisson_solver.jl:141-178,
                                                        src/DistributedComputations/distributed_transpose.jl:25-191
The kernels (pack/unpack, FFT passes, fused tendencies, pressure correction) live in libnhstep.so (csrc/dist.cuh); this
module only places the messages between them.  The index helpers below restate the buffer layouts of csrc/dist.cuh so
that the host logic can be tested on CPU with the ``gloo`` backend (tests/test_distributed_cpu.py).
"""
import ctypes as C
import os

import numpy as np

from . import _lib as L
from ._lib import lib, check


# ---- partition arithmetic (pure host logic) ------------------------------------------------------------------------
class Partition:
    """Partition(x=1, y=R, z=1) or Partition(x=R): slab decompositions (§8e).  The kernels partition in y; an x-slab run is the same
    run on the mirror image of the problem (x ↔ y, u ↔ v, f → −f, west/east ↔ south/north), which the discretisation maps onto
    itself — see DistributedNonhydrostaticModel.  Pencils (x and y together) and z partitions are not built."""

    def __init__(self, x=None, y=None, z=1):
        xi, yi = (1 if x is None else int(x)), (1 if y is None else int(y))
        if z != 1:
            raise NotImplementedError("the vertical direction is not partitioned (the reference's solvers do not either)")
        if xi != 1 and yi != 1:
            raise NotImplementedError("pencil (x × y) decompositions are not built: Partition(x=R) or Partition(y=R)")
        self.x, self.y, self.z = xi, yi, 1
        # Partition(R) / Partition(x=R) slabs along x (also with R = 1, which runs the mirrored code path on one rank); everything
        # else — Partition(), Partition(y=R), Partition(1, R) — along y
        self.direction = "x" if (x is not None and yi == 1 and (xi > 1 or y is None)) else "y"

    @property
    def ranks(self):
        return self.x * self.y


# ---- x-slabs as y-slabs of the mirrored problem ----------------------------------------------------------------------
_SWAP_NAME = {"u": "v", "v": "u"}
_SWAP_SIDE = {"west": "south", "east": "north", "south": "west", "north": "east", "bottom": "bottom", "top": "top"}


def _swap3(t):
    return None if t is None else (t[1], t[0], t[2])


def _mirror_model_arguments(advection, kw):
    """Model keyword arguments of the mirrored (x ↔ y) problem: per-direction advection schemes exchanged, the Coriolis
    parameter negated (a reflection reverses the sense of rotation), boundary conditions moved to the mirrored field and wall."""
    from . import FluxFormAdvection, FPlane
    from .model import FieldBoundaryConditions
    kw = dict(kw)
    if isinstance(advection, FluxFormAdvection):
        advection = FluxFormAdvection(x=advection.y, y=advection.x, z=advection.z)
    if kw.get("coriolis") is not None:
        kw["coriolis"] = FPlane(f=-kw["coriolis"].f)
    if kw.get("boundary_conditions"):
        out = {}
        for name, fbcs in kw["boundary_conditions"].items():
            sides = {}
            for side, bc in fbcs.sides.items():
                if np.ndim(bc.condition) != 0:
                    raise NotImplementedError("array boundary conditions are not supported on an x-slab partition; use Partition(y=R)")
                sides[_SWAP_SIDE[side]] = bc
            out[_SWAP_NAME.get(name, name)] = FieldBoundaryConditions(**sides)
        kw["boundary_conditions"] = out
    return advection, kw


def local_slab(Ny, R, rank):
    """(first global row, zero-based; number of rows) owned by ``rank`` — equal slabs, as Partition(1, R, 1) with Ny % R == 0."""
    if Ny % R:
        raise ValueError(f"Ny = {Ny} is not divisible by the {R} ranks")
    return rank * (Ny // R), Ny // R


def neighbours(rank, R):
    """(previous, next) rank along the periodic y direction (distributed_architectures.jl:182-285, RankConnectivity south/north)."""
    return (rank - 1) % R, (rank + 1) % R


def spectral_block_width(Nx, R):
    """Cx: columns of the half spectrum (kx = 0…Nx/2) owned by one rank, rounded up to the kernels' 8-column tile."""
    cols = Nx // 2 + 1
    return ((cols + R - 1) // R + 7) // 8 * 8


def blocked_index(kx, jl, kz, Cx, NyL, Nz):
    """Offset of spectral element (kx, jl, kz) in the z-local buffer A: R blocks of Cx columns, block d goes to rank d."""
    return (kx % Cx) + Cx * (jl + NyL * (kz + Nz * (kx // Cx)))


def transposed_index(kxl, j, kz, Cx, NyL, Nz):
    """Offset of (local column kxl, GLOBAL row j, kz) in the y-local buffer B after the all-to-all (block s = rows of rank s)."""
    return kxl + Cx * ((j % NyL) + NyL * (kz + Nz * (j // NyL)))


# ---- messages ------------------------------------------------------------------------------------------------------
def exchange_y_halos(send_lo, send_hi, recv_lo, recv_hi, rank, R, group=None):
    """send_lo → previous rank's recv_hi, send_hi → next rank's recv_lo (periodic).  1-D torch tensors of equal length."""
    import torch.distributed as dist
    if R == 1:
        recv_hi.copy_(send_lo)
        recv_lo.copy_(send_hi)
        return
    prev, nxt = neighbours(rank, R)
    if dist.get_backend(group) == "gloo":
        # gloo matches by (peer, tag)
        reqs = [dist.isend(send_lo, prev, group=group, tag=0), dist.isend(send_hi, nxt, group=group, tag=1),
                dist.irecv(recv_hi, nxt, group=group, tag=0), dist.irecv(recv_lo, prev, group=group, tag=1)]
        for r in reqs:
            r.wait()
        return
    # NCCL matches by order per peer: with R = 2 both messages go to the same peer, whose first send is its send_lo
    ops = [dist.P2POp(dist.isend, send_lo, prev, group), dist.P2POp(dist.isend, send_hi, nxt, group),
           dist.P2POp(dist.irecv, recv_hi, nxt, group), dist.P2POp(dist.irecv, recv_lo, prev, group)]
    for r in dist.batch_isend_irecv(ops):
        r.wait()            # stream-ordered for NCCL: the current stream waits, the host does not


def transpose_blocks(dst, src, R, group=None):
    """Block d of ``src`` → rank d, received as block (sender's rank) of ``dst``: transpose_z_to_y! / transpose_y_to_z!."""
    import torch.distributed as dist
    if R == 1:
        dst.copy_(src)
        return
    dist.all_to_all_single(dst, src, group=group)


def _torch_device(index):
    import torch
    return torch.device("cuda", index)


class _DeviceMemory:
    """A raw device allocation of the library, exposed to torch without a copy."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class Distributed:
    """Distributed(partition=Partition(y=R)): rank and size come from the default ``torch.distributed`` group."""

    def __init__(self, partition=None, group=None):
        import torch.distributed as dist
        self.group = group
        if dist.is_available() and dist.is_initialized():
            self.rank, self.size = dist.get_rank(group), dist.get_world_size(group)
        else:
            self.rank, self.size = 0, 1
        self.partition = partition or Partition(y=self.size)
        if self.partition.ranks != self.size:
            raise ValueError(f"Partition(x={self.partition.x}, y={self.partition.y}) does not match the {self.size} ranks of the "
                             "process group")


class DistributedNonhydrostaticModel:
    """NonhydrostaticModel on a (Periodic, Periodic, Periodic) RectilinearGrid split into y-slabs over the ranks of ``arch``.

    ``size``, ``x``, ``y``, ``z`` describe the GLOBAL grid; each rank holds rows ``local_slab(Ny, R, rank)``.

    With ``Partition(x=R)`` (``mirrored``) each rank holds columns ``local_slab(Nx, R, rank)`` instead: internally the model is the
    y-slab model of the problem reflected in the plane x = y, and ``set`` / ``set_local`` / ``local_interior`` / ``gather`` translate
    names (u ↔ v) and array axes, so that a caller only ever sees its own orientation.  ``self.model``, ``self.grid``,
    ``self.velocities`` … are the INTERNAL (mirrored) objects.
    """

    def __init__(self, arch, FT=np.float64, size=None, x=None, y=None, z=None, halo=(3, 3, 3), advection=None,
                 timestepper="RungeKutta3", device=0, p2p=False, topology=None, **kw):
        import torch
        from . import RectilinearGrid, Periodic, WENO
        from .model import NonhydrostaticModel
        self.arch, self.rank, self.R = arch, arch.rank, arch.size
        # Partition(x=R): the run is the y-slab run of the mirrored problem; set / gather / local_interior mirror the arrays back
        self.mirrored = arch.partition.direction == "x"
        if self.mirrored:
            size, halo, topology, x, y = _swap3(size), _swap3(halo), _swap3(topology), y, x
            advection, kw = _mirror_model_arguments(advection, kw)
        self.global_size = tuple(int(n) for n in size)
        Nx, Ny, Nz = self.global_size
        self.j0, self.NyL = local_slab(Ny, self.R, self.rank)
        y0, y1 = float(y[0]), float(y[1])
        ylo = y0 + (y1 - y0) * self.rank / self.R
        yhi = y0 + (y1 - y0) * (self.rank + 1) / self.R
        self.global_grid_spec = dict(FT=FT, size=self.global_size, x=x, y=y, z=z, halo=halo)
        self.grid = RectilinearGrid(FT, size=(Nx, self.NyL, Nz), x=x, y=(ylo, yhi), z=z, halo=halo,
                                    topology=topology or (Periodic, Periodic, Periodic), device=device)
        self.model = NonhydrostaticModel(self.grid, advection=advection or WENO(FT, 5), timestepper=timestepper, **kw)
        self._h = self.model._h
        check(lib().nh_dist_configure(self._h, self.rank, self.R, y0, y1, Ny))
        self.device = _torch_device(device)
        self.FT = np.dtype(FT)
        tdt = torch.float64 if self.FT == np.float64 else torch.float32
        cdt = torch.complex128 if self.FT == np.float64 else torch.complex64
        self._buf = {}
        for which in range(6):
            p, nb = C.c_void_p(), C.c_int64()
            check(lib().nh_dist_buffer(self._h, which, C.byref(p), C.byref(nb)))
            t = torch.as_tensor(_DeviceMemory(p.value, nb.value), device=self.device)
            self._buf[which] = t.view(cdt if which >= L.DIST_SPECTRAL_A else tdt)
        self.velocities, self.pressures = self.model.velocities, self.model.pressures
        self.Cx = spectral_block_width(Nx, self.R)
        self.halo = self.model.halo
        self._uvw = (1 << L.FIELD_U) | (1 << L.FIELD_V) | (1 << L.FIELD_W)
        self.p2p = False
        mode = C.c_int32()
        check(lib().nh_dist_mode(self._h, C.byref(mode)))
        # complete path (anything but the fused triply periodic configuration): the library drives the whole step itself
        # over peer memory, so the peer-to-peer mode is not optional and the host loop is the single-GPU one
        self.generic = mode.value == 2
        self.tracers = self.model.tracers
        if p2p or self.generic:
            self._enable_peer_to_peer()

    def _enable_peer_to_peer(self):
        """Exchange CUDA IPC handles so that the kernels store transposes and halo rows straight into the peers' buffers
        (NVLink peer memory); the messages then shrink to one stream-ordered barrier per piece."""
        import torch
        import torch.distributed as dist
        n = C.c_int64()
        check(lib().nh_dist_ipc_export(self._h, None, 0, C.byref(n)))
        blob = (C.c_ubyte * n.value)()
        check(lib().nh_dist_ipc_export(self._h, blob, n.value, C.byref(n)))
        mine = torch.frombuffer(bytearray(blob), dtype=torch.uint8)
        if self.R > 1:
            allb = torch.empty(self.R * n.value, dtype=torch.uint8, device=self.device)
            dist.all_gather_into_tensor(allb, mine.to(self.device), group=self.arch.group)
            allb = allb.cpu()
        else:
            allb = mine
        raw = (C.c_ubyte * allb.numel()).from_buffer_copy(allb.numpy().tobytes())
        check(lib().nh_dist_ipc_import(self._h, raw, allb.numel()))
        self._flag = torch.zeros(1, device=self.device)
        self._nccl_barrier = bool(os.environ.get("NHSTEP_NCCL_BARRIER"))
        if self.R > 1:
            import torch.distributed as dist
            dist.barrier(group=self.arch.group)          # every rank has opened its peers' flag arrays before the first epoch
        self.p2p = True

    def _barrier(self):
        """All ranks' kernels issued so far have completed before anything issued after it starts (stream-ordered):
        the library's flag barrier over NVLink (nh_dist_barrier), or a one-element NCCL all-reduce with
        NHSTEP_NCCL_BARRIER=1 (the cross-check)."""
        if self.R > 1:
            if self._nccl_barrier:
                import torch.distributed as dist
                dist.all_reduce(self._flag, group=self.arch.group)
            else:
                check(lib().nh_dist_barrier(self._h))

    def transpose(self, direction):
        """transpose_z_to_y! (0) / transpose_y_to_z! (1) between the two spectral buffers (distributed_transpose.jl:25-191):
        peer stores over NVLink bracketed by barriers in the peer-to-peer mode, one all-to-all otherwise."""
        b = self._buf
        if self.p2p:
            self._barrier()
            check(lib().nh_dist_transpose(self._h, int(direction)))
            self._barrier()
        elif direction == 0:
            transpose_blocks(b[L.DIST_SPECTRAL_B], b[L.DIST_SPECTRAL_A], self.R, self.arch.group)
        else:
            transpose_blocks(b[L.DIST_SPECTRAL_A], b[L.DIST_SPECTRAL_B], self.R, self.arch.group)

    # -- pieces ---------------------------------------------------------------------------------------------------
    def exchange(self, mask, width):
        """fill_halo_regions! in y for the fields of ``mask``: pack → NCCL send/recv → unpack."""
        n = C.c_int64()
        check(lib().nh_dist_halo(self._h, mask, width, 0, C.byref(n)))
        b, n = self._buf, n.value
        exchange_y_halos(b[L.DIST_SEND_LO][:n], b[L.DIST_SEND_HI][:n], b[L.DIST_RECV_LO][:n], b[L.DIST_RECV_HI][:n],
                         self.rank, self.R, self.arch.group)
        check(lib().nh_dist_halo(self._h, mask, width, 1, None))

    def pressure_phase(self, stage_dt, trailing_barrier=True):
        """calculate_pressure_correction! + pressure_correct_velocities! + fill_halo_regions!(velocities).  In the peer-to-peer
        mode ``trailing_barrier=False`` leaves the barrier that publishes the neighbours' halo rows to the caller (time_step puts
        the interior part of the next tendency sweep in front of it)."""
        h, b = self._h, self._buf
        if self.p2p:
            # the kernels have already stored into the peers' buffers: only barriers remain
            self._barrier()                                   # v*[Ny+1] written by the next rank's tendency sweep
            check(lib().nh_dist_poisson(h, 0, stage_dt))      # … scatters its z-FFT output into the peers' B
            self._barrier()
            check(lib().nh_dist_poisson(h, 1, stage_dt))      # … scatters the solution back into the peers' A
            self._barrier()
            check(lib().nh_dist_poisson(h, 2, stage_dt))      # … also writes p[0] of the next rank
            self._barrier()
            check(lib().nh_dist_pressure_correct(h, stage_dt))   # … also writes the neighbours' u, v, w halo rows
            if trailing_barrier:
                self._barrier()
            return
        self.exchange(1 << L.FIELD_V, 1)
        check(lib().nh_dist_poisson(h, 0, stage_dt))
        transpose_blocks(b[L.DIST_SPECTRAL_B], b[L.DIST_SPECTRAL_A], self.R, self.arch.group)
        check(lib().nh_dist_poisson(h, 1, stage_dt))
        transpose_blocks(b[L.DIST_SPECTRAL_A], b[L.DIST_SPECTRAL_B], self.R, self.arch.group)
        check(lib().nh_dist_poisson(h, 2, stage_dt))
        self.exchange(1 << L.FIELD_PNHS, 1)
        check(lib().nh_dist_pressure_correct(h, stage_dt))
        self.exchange(self._uvw, self.halo[1])

    def time_step(self, dt, euler=False, n_steps=1):
        """time_step!(model, Δt) — runge_kutta_3.jl:77-151 / quasi_adams_bashforth_2.jl:74-115 on the slab partition."""
        if self.generic:
            check(lib().nh_time_step(self._h, float(dt), int(euler), int(n_steps)))
            return
        stages = 3 if self.model.timestepper_name in ("RungeKutta3", "RK3") else 1
        sdt = C.c_double()
        if self.clock.iteration == 0:
            # update_state! at iteration 0 (runge_kutta_3.jl:81): halos of fields that were set without the set! tail
            check(lib().nh_fill_halo_regions(self._h, self._uvw, 1))
            self.exchange(self._uvw, self.halo[1])
            if self.p2p:
                self._barrier()
        if self.p2p and os.environ.get("NHSTEP_OVERLAP", "0") == "1":
            # (off by default: measured 5.48 vs 5.31 ms per step at 2 GPUs, 256³ per GPU — the second launch of two tile rows
            # costs more than the skew it hides; profiles/r2_scaling.md)
            # interior / boundary overlap (interleave_communication_and_computation.jl:9-67): the tile rows of the sweep that need
            # no neighbour's halo row are issued BEFORE the barrier that closes the previous stage's pressure correction, so that
            # barrier's wait (the skew between the ranks) is spent computing; the two boundary tile rows follow it
            pending = False
            for _ in range(n_steps):
                for s in range(stages):
                    check(lib().nh_dist_tendency_substep_part(self._h, float(dt), s, int(euler), 1, C.byref(sdt)))
                    if pending:
                        self._barrier()
                    check(lib().nh_dist_tendency_substep_part(self._h, float(dt), s, int(euler), 2, C.byref(sdt)))
                    self.pressure_phase(sdt.value, trailing_barrier=False)
                    pending = True
            self._barrier()              # the state is consistent on every rank when time_step returns
            return
        for _ in range(n_steps):
            for s in range(stages):
                check(lib().nh_dist_tendency_substep(self._h, float(dt), s, int(euler), C.byref(sdt)))
                self.pressure_phase(sdt.value)

    def time_step_profiled(self, dt, euler=False):
        """One time_step! with CUDA events between its pieces (ms per step): the three kernel phases of the single-GPU
        profile plus the two kinds of messages (each including its pack/unpack kernels and the NCCL call)."""
        import torch
        h, b, g = self._h, self._buf, self.arch.group
        ev, keys = [torch.cuda.Event(enable_timing=True)], []

        def mark(key):
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            ev.append(e)
            keys.append(key)

        stages = 3 if self.model.timestepper_name in ("RungeKutta3", "RK3") else 1
        sdt = C.c_double()
        ev[0].record()
        for s in range(stages):
            check(lib().nh_dist_tendency_substep(h, float(dt), s, int(euler), C.byref(sdt)))
            mark("tendencies_substep")
            if self.p2p:
                for phase, key in ((0, "pressure_solve"), (1, "pressure_solve"), (2, "pressure_solve"), (3, "pressure_correct")):
                    self._barrier()
                    mark("halo_exchange")               # barrier only: the data moved inside the kernels
                    if phase < 3:
                        check(lib().nh_dist_poisson(h, phase, sdt.value))
                    else:
                        check(lib().nh_dist_pressure_correct(h, sdt.value))
                    mark(key)
                self._barrier()
                mark("halo_exchange")
                continue
            self.exchange(1 << L.FIELD_V, 1)
            mark("halo_exchange")
            check(lib().nh_dist_poisson(h, 0, sdt.value))
            mark("pressure_solve")
            transpose_blocks(b[L.DIST_SPECTRAL_B], b[L.DIST_SPECTRAL_A], self.R, g)
            mark("transpose")
            check(lib().nh_dist_poisson(h, 1, sdt.value))
            mark("pressure_solve")
            transpose_blocks(b[L.DIST_SPECTRAL_A], b[L.DIST_SPECTRAL_B], self.R, g)
            mark("transpose")
            check(lib().nh_dist_poisson(h, 2, sdt.value))
            mark("pressure_solve")
            self.exchange(1 << L.FIELD_PNHS, 1)
            mark("halo_exchange")
            check(lib().nh_dist_pressure_correct(h, sdt.value))
            mark("pressure_correct")
            self.exchange(self._uvw, self.halo[1])
            mark("halo_exchange")
        torch.cuda.synchronize()
        out = {"tendencies_substep": 0.0, "pressure_solve": 0.0, "pressure_correct": 0.0, "halo_exchange": 0.0, "transpose": 0.0}
        for q, key in enumerate(keys):
            out[key] += ev[q].elapsed_time(ev[q + 1])
        return out

    # -- state ----------------------------------------------------------------------------------------------------
    def set(self, enforce_incompressibility=True, **global_arrays):
        """set!(model; u=…, v=…, w=…) from GLOBAL (Nx, Ny, Nz) arrays; every rank keeps its own rows."""
        if self.mirrored:
            global_arrays = {_SWAP_NAME.get(n, n): np.asarray(a).transpose(1, 0, 2) for n, a in global_arrays.items()}
        rows = {}
        for name, a in global_arrays.items():
            a = np.asarray(a)
            if a.shape[:2] != self.global_size[:2]:
                shown = (_SWAP_NAME.get(name, name), _swap3(self.global_size)) if self.mirrored else (name, self.global_size)
                raise ValueError(f"{shown[0]} must have the global size {shown[1]}")
            rows[name] = a[:, self.j0:self.j0 + self.NyL, :]
        self._set_local_internal(enforce_incompressibility, **rows)

    def _field(self, name):
        for ns in (self.velocities, self.tracers, self.pressures):
            if name in ns:
                return ns[name]
        raise KeyError(name)

    def set_local(self, enforce_incompressibility=True, **local_arrays):
        """As ``set`` with this rank's own (Nx, Ny/R, Nz) rows — (Nx/R, Ny, Nz) columns on an x-slab partition."""
        if self.mirrored:
            local_arrays = {_SWAP_NAME.get(n, n): np.asarray(a).transpose(1, 0, 2) for n, a in local_arrays.items()}
        self._set_local_internal(enforce_incompressibility, **local_arrays)

    def _set_local_internal(self, enforce_incompressibility=True, **local_arrays):
        if self.generic:
            # every rank sets the same fields in the same order (each set fills halos, i.e. meets the others at a barrier)
            from .model import set_model
            set_model(self.model, enforce_incompressibility, **{n: local_arrays[n] for n in sorted(local_arrays)})
            return
        for name, a in local_arrays.items():
            self.velocities[name].set(a)
        check(lib().nh_fill_halo_regions(self._h, self._uvw, 1))
        self.exchange(self._uvw, self.halo[1])
        if enforce_incompressibility:
            self.pressure_phase(1.0)

    def local_interior(self, name):
        """This rank's part of a field, in the user's orientation."""
        if self.mirrored:
            return self._field(_SWAP_NAME.get(name, name)).interior().transpose(1, 0, 2)
        return self._field(name).interior()

    def gather(self, name):
        """Global (Nx, Ny, Nz) array of a field on every rank (diagnostics / tests only)."""
        import torch
        import torch.distributed as dist
        internal = _SWAP_NAME.get(name, name) if self.mirrored else name
        loc = np.ascontiguousarray(self._field(internal).interior().transpose(1, 0, 2))       # the partitioned direction first
        if self.R == 1:
            glob = loc.transpose(1, 0, 2)
        else:
            t = torch.from_numpy(loc).to(self.device)
            out = [torch.empty_like(t) for _ in range(self.R)]
            dist.all_gather(out, t, group=self.arch.group)
            glob = torch.cat(out, dim=0).cpu().numpy().transpose(1, 0, 2)
        return glob.transpose(1, 0, 2) if self.mirrored else glob

    def max_abs_divergence(self):
        import torch
        import torch.distributed as dist
        v = self.model.max_abs_divergence()
        if self.R == 1:
            return v
        t = torch.tensor([v], device=self.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.arch.group)
        return float(t.item())

    def synchronize(self):
        self.model.synchronize()

    def launch_count(self):
        return self.model.launch_count()

    @property
    def clock(self):
        return self.model.clock

"""Checkpointer — the reference's checkpoint layout (src/OutputWriters/checkpointer.jl) written as a plain HDF5 file.

What the reference writes with JLD2 (``write_output!(c::Checkpointer, model)``, checkpointer.jl:177-201) and reads back in
``set!(model, filepath)`` (:223-260) and ``set_time_stepper_tendencies!`` (:262-282):

    NonhydrostaticModel/checkpointed_properties
    NonhydrostaticModel/grid                      (a serialized Julia struct)
    NonhydrostaticModel/clock                     (a serialized Julia struct: time, last_Δt, last_stage_Δt, iteration, stage)
    NonhydrostaticModel/timestepper/Gⁿ/<name>/data, …/G⁻/<name>/data
    NonhydrostaticModel/<name>/data               = Array(parent(field)), halos included, for u, v, w and every tracer

Here the same addresses hold the same arrays (``data`` has the parent array's Julia size, see hdf5.py) next to ``location``;
the two Julia structs are written property by property the way the reference's own ``saveproperty!`` writes structs that have to
be readable outside Julia (output_writer_utils.jl:74-85) — ``grid/Nx``, ``grid/Δxᶜᵃᵃ``, ``grid/z/Δᵃᵃᶜ`` … — except the clock, which
is one compound scalar with the struct's members (``time``, ``last_Δt``, ``last_stage_Δt``, ``iteration``, ``stage``).  The file is a classic HDF5 file (any libhdf5 opens it; ``julia/OceananigansB200Ext.jl`` reads it with
HDF5.jl into a Julia model), NOT a JLD2 file: JLD2's encoding of Julia structs cannot be produced, or a Julia-written file
parsed, without Julia, and this image has neither Julia nor libhdf5.  File names follow the reference
(``<prefix>_iteration<N>``) with the extension ``.h5``.

Slab runs: every rank checkpoints its own local model (``Checkpointer(m.model, prefix=f"checkpoint_rank{m.rank}")``, the
``_rank<r>`` naming of the reference's distributed output files, output_writer_utils.jl:220-226) — the parent arrays carry the
neighbours' halo rows, so ``set_from_checkpoint(m.model, path)`` on every rank (without the update_state! of ``pickup``, which would
refill the y halos locally) continues the run bit for bit (tests/test_zw_slab_checkpoint.py).

Only host code lives here: the arrays travel through ``nh_get_field`` / ``nh_set_field`` (parent region) and the clock through
``nh_clock_get`` / ``nh_clock_set``.
"""
import glob
import os
import warnings

import numpy as np

from . import hdf5

ADDRESS = "NonhydrostaticModel"                         # checkpointer_address(::NonhydrostaticModel), checkpointer.jl:121
EXTENSION = ".h5"
_LOCATION = {"c": "Center", "f": "Face"}


def default_checkpointed_properties(model):
    """checkpointer.jl:20-26 (``particles`` is always ``nothing`` on this path and is written as an empty group)."""
    return ["grid", "particles", "clock", "timestepper"]


class Checkpointer:
    """Checkpointer(model; schedule, dir, prefix, overwrite_existing, verbose, cleanup, properties) — checkpointer.jl:76-115.
    ``schedule`` is an iteration interval (IterationInterval(n)) or None; ``maybe_write`` applies it."""

    def __init__(self, model, schedule=None, dir=".", prefix="checkpoint", overwrite_existing=False, verbose=False,
                 cleanup=False, properties=None):
        properties = list(default_checkpointed_properties(model) if properties is None else properties)
        required = ["grid", "particles", "clock"]
        if model.timestepper_name in ("QuasiAdamsBashforth2", "AB2"):
            required.append("timestepper")
        for rp in required:
            if rp not in properties:
                warnings.warn(f"{rp} is required for checkpointing. It will be added to checkpointed properties")
                properties.append(rp)
        for p in properties:
            if not isinstance(p, str):
                raise ValueError(f"Property {p} to be checkpointed must be a string (a Symbol in the reference).")
            if p not in ("grid", "particles", "clock", "timestepper"):
                raise ValueError(f"Cannot checkpoint {p}, it is not a model property!")
        os.makedirs(dir, exist_ok=True)
        self.schedule, self.dir, self.prefix, self.properties = schedule, dir, prefix, properties
        self.overwrite_existing, self.verbose, self.cleanup = overwrite_existing, verbose, cleanup

    def maybe_write(self, model):
        """The schedule check the reference's Simulation loop performs before write_output!."""
        if self.schedule and model.clock.iteration % int(self.schedule) == 0:
            return write_output(self, model)
        return None


def checkpoint_superprefix(prefix):
    return prefix + "_iteration"                         # checkpointer.jl:124


def checkpoint_path(which, checkpointer):
    """checkpoint_path(iteration::Int, c) / (filepath::String, …) / (pickup::Bool, c) — checkpointer.jl:134-164."""
    if isinstance(which, (bool, np.bool_)):
        paths = _checkpoint_files(checkpointer)
        if not paths:
            warnings.warn("pickup=true but no checkpoints were found. Simulation will run without picking up.")
            return None
        return latest_checkpoint(checkpointer, paths)
    if isinstance(which, (str, os.PathLike)):
        return os.fspath(which)
    return os.path.join(checkpointer.dir, f"{checkpoint_superprefix(checkpointer.prefix)}{int(which)}{EXTENSION}")


def _checkpoint_files(checkpointer):
    pattern = glob.escape(os.path.join(checkpointer.dir, checkpoint_superprefix(checkpointer.prefix))) + "*" + EXTENSION
    return sorted(glob.glob(pattern))


def latest_checkpoint(checkpointer, filepaths):
    """The file with the largest iteration in its name — checkpointer.jl:166-173."""
    lead = len(checkpoint_superprefix(checkpointer.prefix))
    iterations = [int(os.path.basename(p)[lead:-len(EXTENSION)]) for p in filepaths]
    return filepaths[int(np.argmax(iterations))]


def cleanup_checkpoints(checkpointer):
    """checkpointer.jl:203-208: keep only the latest checkpoint."""
    paths = _checkpoint_files(checkpointer)
    if not paths:
        return
    keep = latest_checkpoint(checkpointer, paths)
    for p in paths:
        if p != keep:
            os.remove(p)


# ---------------------------------------------------------------------------------------------------------------------
# model ⇄ tree
# ---------------------------------------------------------------------------------------------------------------------

def _field_group(field):
    """serializeproperty!(file, address, f::Field) — output_writer_utils.jl:144-149 (location and data; the indices are the
    whole field and the boundary conditions live in the model description on this path)."""
    return {"location": np.array([_LOCATION[c] for c in field.loc], dtype="S6"), "data": field.parent()}


def grid_tree(model):
    """The RectilinearGrid struct (rectilinear_grid.jl:1-23) property by property, with the inflated halo of the model's grid.
    The node and z-spacing arrays hold the interior range (the reference's OffsetArrays also carry the halo nodes)."""
    g = model.grid
    FT = g.FT
    dx, dy, dzc, dzf = g.spacings()
    N, H = g.N, model.halo
    ext = g.extents()
    own_Hz = (len(dzc) - N[2]) // 2                                    # spacings() carries the halo of the grid object itself
    tree = {
        "Nx": np.int64(N[0]), "Ny": np.int64(N[1]), "Nz": np.int64(N[2]),
        "Hx": np.int64(H[0]), "Hy": np.int64(H[1]), "Hz": np.int64(H[2]),
        "Lx": ext[0], "Ly": ext[1], "Lz": ext[2],
        "topology": np.array(list(g.topology), dtype="S8"),
        "Δxᶠᵃᵃ": dx, "Δxᶜᵃᵃ": dx, "Δyᵃᶠᵃ": dy, "Δyᵃᶜᵃ": dy,
        "xᶠᵃᵃ": g.nodes(0, "f"), "xᶜᵃᵃ": g.nodes(0, "c"), "yᵃᶠᵃ": g.nodes(1, "f"), "yᵃᶜᵃ": g.nodes(1, "c"),
        "z": {"cᵃᵃᶠ": g.nodes(2, "f"), "cᵃᵃᶜ": g.nodes(2, "c")},
    }
    if g.z_stretched:
        tree["z"]["Δᵃᵃᶜ"] = np.asarray(dzc[own_Hz:own_Hz + N[2]])
        tree["z"]["Δᵃᵃᶠ"] = np.asarray(dzf[own_Hz:own_Hz + N[2] + 1])
    else:
        tree["z"]["Δᵃᵃᶜ"] = FT(dzc[own_Hz]) if len(dzc) else FT(1)
        tree["z"]["Δᵃᵃᶠ"] = tree["z"]["Δᵃᵃᶜ"]
    return tree


CLOCK_DTYPE = np.dtype([("time", "<f8"), ("last_Δt", "<f8"), ("last_stage_Δt", "<f8"), ("iteration", "<i8"), ("stage", "<i8")])


def clock_tree(clock):
    """Clock (clock.jl:16-22) as ONE compound scalar with the struct's member names and order — the HDF5 form JLD2 gives a Julia
    struct, so that `file["…/clock"].iteration` / `.time` / `.last_Δt` (checkpointer.jl:251-256) resolve on whatever a reader builds
    from a compound (JLD2: a reconstructed struct, HDF5.jl: a NamedTuple, h5py: a structured scalar)."""
    return np.array((clock.time, clock.last_dt, clock.last_stage_dt, clock.iteration, clock.stage), dtype=CLOCK_DTYPE)


def state_tree(model, properties=None):
    """Everything write_output! puts in the file (checkpointer.jl:184-193) as a nested dict."""
    properties = default_checkpointed_properties(model) if properties is None else properties
    body = {"checkpointed_properties": np.array(list(properties), dtype="S")}
    fields = model.prognostic_fields()
    if "grid" in properties:
        body["grid"] = grid_tree(model)
    if "clock" in properties:
        body["clock"] = clock_tree(model.clock)
    if "particles" in properties:
        body["particles"] = {}
    if "timestepper" in properties:
        body["timestepper"] = {"name": model.timestepper_name,
                               "Gⁿ": {f.name: _field_group(model.timestepper.Gn[f.name]) for f in fields},
                               "G⁻": {f.name: _field_group(model.timestepper.Gm[f.name]) for f in fields}}
    for f in fields:
        body[f.name] = _field_group(f)
    return {ADDRESS: body}


def write_output(checkpointer, model):
    """write_output!(c::Checkpointer, model) — checkpointer.jl:177-201.  Returns the file path."""
    path = checkpoint_path(int(model.clock.iteration), checkpointer)
    if os.path.exists(path) and not checkpointer.overwrite_existing:
        raise FileExistsError(f"{path} exists and overwrite_existing is false")
    hdf5.write(path, state_tree(model, checkpointer.properties))
    if checkpointer.verbose:
        print(f"Checkpointing done: {path}, size {os.path.getsize(path)} bytes")
    if checkpointer.cleanup:
        cleanup_checkpoints(checkpointer)
    return path


def _same_grid(model, G):
    g = model.grid
    want = grid_tree(model)
    try:
        same = all(int(G.read(k)) == int(want[k]) for k in ("Nx", "Ny", "Nz", "Hx", "Hy", "Hz"))
        same = same and list(G.read("topology")) == list(g.topology)
        for k in ("Δxᶜᵃᵃ", "Δyᵃᶜᵃ"):
            same = same and np.array_equal(np.asarray(G.read(k)), np.asarray(want[k]))
        same = same and np.array_equal(np.asarray(G.read("z/Δᵃᵃᶜ")), np.asarray(want["z"]["Δᵃᵃᶜ"]))
    except KeyError:
        return False
    return bool(same)


def set_from_checkpoint(model, filepath):
    """set!(model, filepath::AbstractString) — checkpointer.jl:223-260 with set_time_stepper_tendencies! (:262-282): parent data of
    every prognostic field found in the file, the tendencies Gⁿ and G⁻ where the file has them (the reference restores them for
    QuasiAdamsBashforth2 only; for RungeKutta3 they are recomputed by the update_state! of ``pickup``), and the clock's iteration,
    time and last_Δt (+ stage and last_stage_Δt, which the reference's struct also carries)."""
    f = hdf5.File(filepath)
    if ADDRESS not in f:
        raise KeyError(f"{filepath}: no {ADDRESS} group")
    A = f[ADDRESS]
    try:
        G = A["grid"] if "grid" in A else None
        same = isinstance(G, hdf5.Group) and _same_grid(model, G)     # a Julia-serialized grid struct cannot be compared here
    except (NotImplementedError, ValueError):
        same = False
    if not same:
        warnings.warn(f"The grid associated with {filepath} and model.grid are not the same!")
    names = A.keys()
    for fld in model.prognostic_fields():
        if fld.name in names:
            fld.set_parent(_checked(A.read(f"{fld.name}/data"), fld, filepath))
        else:
            warnings.warn(f"Field {fld.name} does not exist in checkpoint and could not be restored.")
    if "timestepper" in names:
        T = A["timestepper"]
        for fld in model.prognostic_fields():
            if "Gⁿ" in T and fld.name in T["Gⁿ"].keys():
                G = model.timestepper.Gn[fld.name]
                G.set_parent(_checked(T.read(f"Gⁿ/{fld.name}/data"), G, filepath))
                G = model.timestepper.Gm[fld.name]
                G.set_parent(_checked(T.read(f"G⁻/{fld.name}/data"), G, filepath))
            else:
                warnings.warn(f"Tendencies for {fld.name} do not exist in checkpoint and could not be restored.")
    c = A.read("clock")
    now = model.clock
    model.set_clock(time=float(c["time"]), iteration=int(c["iteration"]), last_dt=float(c["last_Δt"]),
                    stage=int(c.get("stage", now.stage)), last_stage_dt=float(c.get("last_stage_Δt", now.last_stage_dt)))


def _checked(data, field, filepath):
    want = field.parent_size
    if tuple(data.shape) != tuple(want):
        raise ValueError(f"{filepath}: {field.name}/data has size {tuple(data.shape)}, the model's parent array is {tuple(want)}")
    return data


def pickup(model, which, checkpointer=None):
    """run!(simulation; pickup) up to the first time step (Simulations/run.jl:59-62,92-98,203-211): resolve ``which`` (True = latest
    checkpoint of ``checkpointer``, an iteration number, or a file path), ``set!(model, filepath)``, then the ``update_state!`` of
    ``initialize!(sim)``, which fills the halos and recomputes the auxiliary fields and Gⁿ from the restored state."""
    path = checkpoint_path(which, checkpointer)
    if path is None:
        return None
    set_from_checkpoint(model, path)
    model.update_state(compute_tendencies=True)
    return path


# ---------------------------------------------------------------------------------------------------------------------
# the layout of the reference's golden regression files
# ---------------------------------------------------------------------------------------------------------------------

def regression_tree(model):
    """The older checkpoint layout the reference's downloaded regression files have (what get_fields_from_checkpoint reads,
    test/test_nonhydrostatic_regression.jl:7-42): ``velocities/<u|v|w>/data``, ``tracers/<name>/data`` and
    ``timestepper/Gⁿ|G⁻/<name>/data`` at the root of the file, parent arrays."""
    fields = model.prognostic_fields()
    return {"velocities": {f.name: {"data": f.parent()} for f in fields[:3]},
            "tracers": {f.name: {"data": f.parent()} for f in fields[3:]},
            "timestepper": {"Gⁿ": {f.name: {"data": model.timestepper.Gn[f.name].parent()} for f in fields},
                            "G⁻": {f.name: {"data": model.timestepper.Gm[f.name].parent()} for f in fields}}}


def get_fields_from_checkpoint(filename):
    """get_fields_from_checkpoint (test/test_nonhydrostatic_regression.jl:7-42): (solution, Gⁿ, G⁻) dicts of parent arrays by field
    name from a file in the regression layout — the reader a golden-file comparison starts from (the files themselves are remote
    DataDeps, test/data_dependencies.jl:17-38, and JLD2-written: see hdf5.py on what of that container is verified)."""
    f = hdf5.File(filename)
    tracers = f["tracers"].keys() if "tracers" in f else []
    names = ["u", "v", "w"] + list(tracers)
    where = {n: ("velocities" if n in "uvw" else "tracers") for n in names}
    solution = {n: f.read(f"{where[n]}/{n}/data") for n in names}
    Gn = {n: f.read(f"timestepper/Gⁿ/{n}/data") for n in names}
    Gm = {n: f.read(f"timestepper/G⁻/{n}/data") for n in names}
    return solution, Gn, Gm

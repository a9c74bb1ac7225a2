"""Checkpoint files: the HDF5 subset of oceananigans.jl_b200/hdf5.py and the Checkpointer host logic.

CPU part.  The reader is pinned on the one libhdf5-written file in this image (scipy's MATLAB v7.3 test file), the writer on the
bytes libhdf5 put into that file for the same datatype / dataspace / heap / B-tree, and on round trips through the reader.  The
checkpointer's own logic (what is written, what set!(model, filepath) restores, pickup, cleanup) runs here over the ORACLE model
behind an adapter with the host mirror's interface — the reference's test_checkpointer.jl:60-79,98-178,180-194 restated with
bit equality in place of ``≈``.  The GPU version (the library's model through the C ABI) is tests/test_zy_checkpoint_gpu.py.
"""
import os
import struct
import warnings

import numpy as np
import pytest

import oracle as o
from oracle import model as om
import oceananigans_jl_b200.hdf5 as h5
import oceananigans_jl_b200.checkpointer as ck

P, B = "Periodic", "Bounded"


def _libhdf5_file():
    try:
        import scipy.io.matlab.tests as t
    except Exception:
        return None
    p = os.path.join(os.path.dirname(t.__file__), "data", "testhdf5_7.4_GLNX86.mat")
    return p if os.path.exists(p) else None


needs_sample = pytest.mark.skipif(_libhdf5_file() is None, reason="scipy's HDF5 test file is not installed")


# ---- the reader against a file libhdf5 wrote ------------------------------------------------------------------------------

@needs_sample
def test_reader_parses_a_libhdf5_written_file():
    f = h5.File(_libhdf5_file())
    assert f.base == 512 and (f.leaf_k, f.internal_k) == (4, 16)          # MATLAB's 512-byte user block, default tree widths
    assert f.keys() == ["testdouble"]
    d = f["testdouble"]
    assert d.attrs == {"MATLAB_class": "double"}
    a = d.read()
    # MATLAB's 1×9 testdouble = 0:pi/4:2*pi (the same variable scipy's other testdouble_*.mat fixtures hold)
    assert a.shape == (1, 9) and a.dtype == np.float64
    assert np.array_equal(a.ravel(), np.arange(9) * (np.pi / 4))


@needs_sample
def test_writer_emits_the_bytes_libhdf5_emits(tmp_path):
    ref = h5.File(_libhdf5_file())
    msgs = {m[0]: m[2] for m in ref["testdouble"]._messages}
    assert h5._encode_datatype(np.float64) == msgs[h5.MSG_DATATYPE][:20]              # IEEE binary64, little-endian
    assert h5._encode_dataspace((9, 1)) == msgs[h5.MSG_DATASPACE]
    # same one-member group written by us: heap, B-tree node and symbol-table node field for field
    path = tmp_path / "one.h5"
    h5.write(path, {"testdouble": np.asfortranarray((np.arange(9) * (np.pi / 4)).reshape(1, 9))})
    mine, theirs = h5.File(path), ref

    def structures(f):
        b = f.buf
        btree, heap = struct.unpack_from("<QQ", next(m[2] for m in f._messages if m[0] == h5.MSG_SYMBOL_TABLE))
        hp, bt = f.base + heap, f.base + btree
        seg_size, free_at, seg = struct.unpack_from("<QQQ", b, hp + 8)
        seg += f.base
        names = bytes(b[seg:seg + free_at])
        free_block = struct.unpack_from("<QQ", b, seg + free_at)
        tree_head = bytes(b[bt:bt + 24])
        key0, child, key1 = struct.unpack_from("<QQQ", b, bt + 24)
        sn = f.base + child
        snod_head = bytes(b[sn:sn + 8])
        name_off = struct.unpack_from("<Q", b, sn + 8)[0]
        return dict(heap_sig=bytes(b[hp:hp + 8]), names=names, free_next=free_block[0], free_size=seg_size - free_at == free_block[1],
                    tree_head=tree_head, keys=(key0, key1), snod_head=snod_head, name_off=name_off)

    assert structures(mine) == structures(theirs)
    assert np.array_equal(mine.read("testdouble"), theirs.read("testdouble"))
    # object header prefix: version 1, reference count 1, 16-byte prefix
    for f, root in ((mine, 96), (theirs, None)):
        at = f.base + struct.unpack_from("<Q", f.buf, (0 if f is mine else 512) + 24 + 32 + 8)[0]
        version, _r, nmsg, refs, size = struct.unpack_from("<BBHII", f.buf, at)
        assert (version, refs) == (1, 1) and size % 8 == 0 and nmsg >= 1


def _equal(a, b):
    if isinstance(a, dict):
        return isinstance(b, dict) and list(sorted(a)) == list(sorted(b)) and all(_equal(a[k], b[k]) for k in a)
    if isinstance(a, str):
        return a == b
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == "S":
        return [v.decode() for v in a.ravel()] == list(b.ravel())
    return a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b)


def test_write_read_round_trip(tmp_path):
    rng = np.random.default_rng(1)
    tree = {"NonhydrostaticModel": {
        "u": {"data": np.asfortranarray(rng.random((7, 5, 3))), "location": np.array(["Face", "Center", "Center"], dtype="S6")},
        "T": {"data": rng.random((4, 1, 6)).astype(np.float32)},
        "clock": {"time": np.float64(0.1) * 3, "iteration": np.int64(2 ** 40 + 7), "last_Δt": np.float64(np.inf), "stage": np.int64(1)},
        "timestepper": {"Gⁿ": {"u": {"data": rng.random((2, 2, 2))}}, "G⁻": {}, "name": "QuasiAdamsBashforth2"},
        "particles": {},
        **{f"member{i:02d}": np.int32(i) for i in range(21)},              # > 8 links: several symbol-table nodes
        "bytes": np.arange(5, dtype=np.uint8), "flag": True, "i16": np.array([[1, -2], [3, -4]], dtype=np.int16),
    }}
    path = tmp_path / "t.h5"
    h5.write(path, tree)
    f = h5.File(path)
    got = f.tree()
    tree["NonhydrostaticModel"]["flag"] = np.uint8(1)
    assert _equal(tree, got)
    assert f["NonhydrostaticModel"].keys() == sorted(tree["NonhydrostaticModel"], key=lambda s: s.encode())
    assert f.read("NonhydrostaticModel/u/data").flags.f_contiguous
    assert "NonhydrostaticModel/timestepper/Gⁿ/u" in f and "NonhydrostaticModel/v" not in f
    with pytest.raises(KeyError):
        f.read("NonhydrostaticModel/u")
    # dimensions are stored reversed (Julia / Fortran order over the same bytes)
    msgs = {m[0]: m[2] for m in f["NonhydrostaticModel/u/data"]._messages}
    assert h5._decode_dataspace(msgs[h5.MSG_DATASPACE]) == (3, 5, 7)
    with pytest.raises(TypeError):
        h5.write(tmp_path / "bad.h5", {"z": np.zeros(2, dtype=complex)})
    # a compound scalar (the form a Julia struct takes in JLD2): members by name, UTF-8 names, mixed member types
    dt = np.dtype([("time", "<f8"), ("last_Δt", "<f4"), ("iteration", "<i8"), ("stage", "<i2")])
    h5.write(tmp_path / "c.h5", {"clock": np.array((0.3, 0.5, 2 ** 40, -3), dtype=dt)})
    got = h5.File(tmp_path / "c.h5").read("clock")
    assert got == {"time": 0.3, "last_Δt": 0.5, "iteration": 2 ** 40, "stage": -3}
    decoded, used = h5._decode_datatype(h5._encode_datatype(dt))
    assert decoded == dt and used == len(h5._encode_datatype(dt))
    with pytest.raises(ValueError):
        h5.File(__file__)


# ---- the Checkpointer over the oracle model (adapter with the host mirror's interface) --------------------------------------

class _FieldView:
    def __init__(self, f, name):
        self.f, self.name, self.loc = f, name, f.loc

    def parent(self):
        return np.asfortranarray(self.f.data.copy())

    def set_parent(self, a):
        self.f.data[...] = a

    @property
    def parent_size(self):
        return self.f.data.shape


class _GridView:
    def __init__(self, g):
        self.g, self.FT, self.N, self.topology = g, g.FT, g.N, g.topology
        self.z_stretched = not g.regular_dim(2)

    def spacings(self):
        cz = self.g.coords[2]
        n = self.g.Nz + 2 * self.g.Hz
        if self.z_stretched:
            dzc, dzf = cz.dc.r(1 - self.g.Hz, self.g.Nz + self.g.Hz), cz.df.r(1 - self.g.Hz, self.g.Nz + self.g.Hz)
        else:
            dzc = dzf = np.full(n, cz.dc)
        return self.g.coords[0].dc, self.g.coords[1].dc, np.asarray(dzc), np.asarray(dzf)

    def extents(self):
        return self.g.Lx, self.g.Ly, self.g.Lz

    def nodes(self, d, loc):
        c = self.g.coords[d]
        if c.F is None:
            return np.ones(self.N[d], dtype=self.FT)
        n = self.N[d] + (1 if loc == "f" and self.topology[d] == B else 0)
        return (c.F if loc == "f" else c.C).r(1, n)


class _Namespace(dict):
    __getattr__ = dict.__getitem__


class _ModelView:
    """The oracle's NonhydrostaticModel behind the interface checkpointer.py uses of the host mirror's model."""

    def __init__(self, m):
        self.m, self.grid, self.halo = m, _GridView(m.grid), tuple(m.grid.H)
        self.timestepper_name = m.timestepper
        names = m.prognostic_names()
        self.timestepper = _Namespace(Gn=_Namespace((n, _FieldView(m.Gn[n], n)) for n in names),
                                      Gm=_Namespace((n, _FieldView(m.Gm[n], n)) for n in names))

    clock = property(lambda self: self.m.clock)

    def prognostic_fields(self):
        return [_FieldView(f, n) for n, f in zip(self.m.prognostic_names(), self.m.prognostic_fields())]

    def set_clock(self, **values):
        for k, v in values.items():
            setattr(self.m.clock, k, v)

    def update_state(self, compute_tendencies=True):
        om.update_state(self.m, compute_tendencies=compute_tendencies)


def _run(view, dt, stop_iteration, checkpointer=None, pickup=False, write=True):
    """run!(simulation; pickup) reduced to what touches the model (Simulations/run.jl:92-98,203-211 and the writer calls)."""
    if pickup is not False:
        ck.pickup(view, pickup, checkpointer)
    else:
        view.update_state()
    write = write and checkpointer is not None
    if view.clock.iteration == 0 and write:
        checkpointer.maybe_write(view)
    while view.clock.iteration < stop_iteration:
        om.time_step(view.m, dt)
        if write:
            checkpointer.maybe_write(view)


def _thermal_bubble(timestepper):
    """test_checkpointer.jl:60-79."""
    N = 16
    g = o.RectilinearGrid(np.float64, size=(N, N, N), extent=(100, 100, 100))
    m = om.NonhydrostaticModel(g, closure=om.ScalarDiffusivity(nu=4e-2, kappa=4e-2), buoyancy=om.SeawaterBuoyancy(),
                               tracers=("T", "S"), timestepper=timestepper)
    T = np.zeros((N, N, N))
    i1, i2 = round(N / 4), round(3 * N / 4)
    T[i1 - 1:i2, i1 - 1:i2, i1 - 1:i2] += 0.01
    om.set_model(m, enforce_incompressibility=False, T=T)
    return m


def _assert_models_equal(a, b):
    """test_model_equality (test_checkpointer.jl:9-29) with == in place of ≈."""
    for fa, fb in zip(a.prognostic_fields(), b.prognostic_fields()):
        assert np.array_equal(fa.data, fb.data), fa.name
    for n in a.Gn:
        assert np.array_equal(a.Gn[n].data, b.Gn[n].data), "Gn " + n
        assert np.array_equal(a.Gm[n].data, b.Gm[n].data), "Gm " + n


@pytest.mark.parametrize("timestepper", ["RungeKutta3", "QuasiAdamsBashforth2"])
def test_thermal_bubble_checkpointer_output(timestepper, tmp_path):
    """run_checkpointer_tests (test_checkpointer.jl:98-178): checkpoint at iterations 0 and 5, run on to 9; a second model set!
    from the file equals the checkpointed one, and picked up from the path, from `true`, from 0 and from 5 it reaches iteration 9
    in the same state as the uninterrupted run."""
    import copy
    dt = 6.0
    true_model, test_model = _ModelView(_thermal_bubble(timestepper)), _ModelView(_thermal_bubble(timestepper))
    c = ck.Checkpointer(true_model, schedule=5, dir=str(tmp_path), overwrite_existing=True)
    _run(true_model, dt, 5, c)
    checkpointed = copy.deepcopy(true_model.m)
    _run(true_model, dt, 9, c)
    five, zero = str(tmp_path / "checkpoint_iteration5.h5"), str(tmp_path / "checkpoint_iteration0.h5")
    assert os.path.exists(five) and os.path.exists(zero)
    assert ck.checkpoint_path(5, c) == five

    ck.set_from_checkpoint(test_model, five)
    assert test_model.clock.iteration == checkpointed.clock.iteration == 5
    assert test_model.clock.time == checkpointed.clock.time
    assert test_model.clock.last_dt == checkpointed.clock.last_dt
    _assert_models_equal(test_model.m, checkpointed)

    c2 = ck.Checkpointer(test_model, schedule=5, dir=str(tmp_path), overwrite_existing=True)
    for pick in (zero, five, True, 0, 5):
        _run(test_model, dt, 9, c2, pickup=pick, write=False)
        assert test_model.clock.iteration == true_model.clock.iteration == 9
        assert test_model.clock.time == true_model.clock.time
        _assert_models_equal(test_model.m, true_model.m)
    assert ck.checkpoint_path(True, c2) == five                          # the latest one


def test_checkpoint_file_has_the_reference_addresses(tmp_path):
    """The addresses set!(model, filepath) reads (checkpointer.jl:229-256,262-282) and the regression tests read back
    (test_nonhydrostatic_regression.jl:13-32), with parent-sized arrays."""
    view = _ModelView(_thermal_bubble("QuasiAdamsBashforth2"))
    _run(view, 6.0, 2)
    path = ck.write_output(ck.Checkpointer(view, dir=str(tmp_path)), view)
    assert os.path.basename(path) == "checkpoint_iteration2.h5"
    f = h5.File(path)
    A = f["NonhydrostaticModel"]
    assert list(A.read("checkpointed_properties")) == ["grid", "particles", "clock", "timestepper"]
    for name, f_ in zip("uvwTS", view.m.prognostic_fields()):
        for addr in (f"{name}/data", f"timestepper/Gⁿ/{name}/data", f"timestepper/G⁻/{name}/data"):
            assert A.read(addr).shape == f_.data.shape == (22, 22, 22 + (name == "w")), addr
        assert np.array_equal(A.read(f"{name}/data"), f_.data)
    assert list(A.read("w/location")) == ["Center", "Center", "Face"]
    assert A.read("clock") == {"time": 12.0, "last_Δt": 6.0, "last_stage_Δt": 6.0, "iteration": 2, "stage": 1}
    G = A["grid"]
    assert [int(G.read(k)) for k in ("Nx", "Ny", "Nz", "Hx", "Hy", "Hz")] == [16, 16, 16, 3, 3, 3]
    assert G.read("Lz") == 100.0 and G.read("Δxᶜᵃᵃ") == 6.25 and G.read("z/Δᵃᵃᶜ") == 6.25
    assert list(G.read("topology")) == [P, P, B] and G.read("z/cᵃᵃᶠ").shape == (17,)
    with pytest.raises(FileExistsError):
        ck.write_output(ck.Checkpointer(view, dir=str(tmp_path)), view)


def test_checkpointer_cleanup(tmp_path):
    """run_checkpointer_cleanup_tests (test_checkpointer.jl:180-194): IterationInterval(3) with cleanup over 10 iterations leaves
    checkpoint_iteration9 only."""
    g = o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1))
    view = _ModelView(om.NonhydrostaticModel(g, buoyancy=om.SeawaterBuoyancy(), tracers=("T", "S")))
    c = ck.Checkpointer(view, schedule=3, cleanup=True, dir=str(tmp_path))
    _run(view, 0.2, 10, c)
    assert sorted(os.listdir(tmp_path)) == ["checkpoint_iteration9.h5"]


def test_latest_checkpoint_orders_by_iteration_number_and_pickup_without_files_warns(tmp_path):
    view = _ModelView(_thermal_bubble("RungeKutta3"))
    c = ck.Checkpointer(view, dir=str(tmp_path), prefix="run[1]")
    with pytest.warns(UserWarning, match="no checkpoints were found"):
        assert ck.checkpoint_path(True, c) is None
    for it in (200, 1000, 30):
        open(ck.checkpoint_path(it, c), "wb").close()
    assert os.path.basename(ck.checkpoint_path(True, c)) == "run[1]_iteration1000.h5"    # checkpointer.jl:166-173
    assert ck.checkpoint_path("elsewhere/x.h5", c) == "elsewhere/x.h5"


def test_constructor_checks_follow_the_reference(tmp_path):
    """checkpointer.jl:85-108."""
    view = _ModelView(_thermal_bubble("QuasiAdamsBashforth2"))
    with pytest.warns(UserWarning, match="timestepper is required"):
        c = ck.Checkpointer(view, dir=str(tmp_path), properties=["grid", "particles", "clock"])
    assert "timestepper" in c.properties
    with pytest.raises(ValueError, match="not a model property"):
        ck.Checkpointer(view, dir=str(tmp_path), properties=["grid", "particles", "clock", "timestepper", "nonsense"])
    with pytest.raises(ValueError, match="must be a string"):
        ck.Checkpointer(view, dir=str(tmp_path), properties=["grid", "particles", "clock", "timestepper", 3])
    assert os.path.isdir(ck.Checkpointer(view, dir=str(tmp_path / "new" / "dir")).dir)   # mkpath(dir)


def test_set_from_checkpoint_warnings_and_errors(tmp_path):
    view = _ModelView(_thermal_bubble("RungeKutta3"))
    path = ck.write_output(ck.Checkpointer(view, dir=str(tmp_path), properties=["grid", "particles", "clock"]), view)
    # another grid: the reference warns and copies anyway; a parent array of another size cannot be copied
    g = o.RectilinearGrid(np.float64, size=(16, 16, 16), extent=(100, 100, 50))
    other = _ModelView(om.NonhydrostaticModel(g, buoyancy=om.SeawaterBuoyancy(), tracers=("T", "S", "c")))
    with pytest.warns(UserWarning) as rec:
        ck.set_from_checkpoint(other, path)
    text = " | ".join(str(w.message) for w in rec)
    assert "are not the same" in text and "Field c does not exist in checkpoint" in text
    small = _ModelView(om.NonhydrostaticModel(o.RectilinearGrid(np.float64, size=(8, 8, 8), extent=(1, 1, 1))))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with pytest.raises(ValueError, match="parent array"):
            ck.set_from_checkpoint(small, path)


def test_the_gpu_pickup_check_itself_over_the_oracle(tmp_path):
    """tests/test_zy_checkpoint_gpu.py's driver (_check_pickup) run here with the oracle model standing in for the library, so
    that its own logic is exercised on the CPU too."""
    import types
    from test_zy_checkpoint_gpu import _check_pickup

    class PressureView:
        def __init__(self, m):
            self.m = m

        def interior(self):
            return self.m.pNHS.interior.copy()

    def make():
        v = _ModelView(_thermal_bubble("QuasiAdamsBashforth2"))
        v.pressures = _Namespace(pNHS=PressureView(v.m))
        v.update_state()
        return v

    def time_step(v, dt, n_steps=1):
        for _ in range(n_steps):
            om.time_step(v.m, dt)

    fake = types.SimpleNamespace(Checkpointer=ck.Checkpointer, pickup=ck.pickup, time_step=time_step,
                                 set_model=lambda v, path: ck.set_from_checkpoint(v, path))
    _check_pickup(fake, make, 6.0, tmp_path)


# ---- the version-2 structures (JLD2's container): self-consistency only ---------------------------------------------------------

def _v2_file(path, u, clock_dtype, clock):
    """Assemble a file the way JLD2 lays one out — 512-byte text header, superblock 2 with base address 512, OHDR object headers with
    link messages, an OCHK continuation chunk, a committed (shared) compound datatype, layout version 4 — from the HDF5 file-format
    specification.  It is NOT a JLD2-written file: it only checks that the reader follows the specification the way this emitter
    reads it; no checksum is computed (the reader ignores them)."""
    buf = bytearray(b"HDF5-based Julia Data Format, version 0.1.1 (test stand-in)".ljust(512, b"\x00"))
    base = 512

    def alloc(data):
        at = len(buf)
        buf.extend(data)
        return at - base                                               # addresses are relative to the base address

    def msg(mtype, data, flags=0, order=None):
        return struct.pack("<BHB", mtype, len(data), flags) + (struct.pack("<H", order) if order is not None else b"") + data

    def ohdr(messages, flags=0, extra=b""):
        body = b"".join(messages)
        width = 1 << (flags & 3)
        return alloc(b"OHDR" + bytes([2, flags]) + extra + len(body).to_bytes(width, "little") + body + b"\x00" * 4)

    def link(name, addr, wide=False):
        raw = name.encode("utf-8")
        if wide:                                                       # link type + creation order + charset present, 2-byte length
            return struct.pack("<BBBQBH", 1, 0x08 | 0x04 | 0x10 | 1, 0, 7, 1, len(raw)) + raw + struct.pack("<Q", addr)
        return struct.pack("<BBB", 1, 0, len(raw)) + raw + struct.pack("<Q", addr)

    superblock_at = alloc(bytes(48))
    # dataset u/data: dataspace v2 (dimensions reversed), float datatype version 3, fill value v3, layout v4 contiguous
    raw = u.tobytes(order="F")
    data_at = alloc(raw)
    float_t = bytes([0x30 | 1]) + h5._encode_datatype(u.dtype)[1:]
    space = struct.pack("<BBBB", 2, u.ndim, 0, 1) + b"".join(struct.pack("<Q", n) for n in u.shape[::-1])
    dset = ohdr([msg(0x1, space), msg(0x3, float_t, 1), msg(0x5, bytes([3, 0x09])), msg(0x8, struct.pack("<BBQQ", 4, 1, data_at, len(raw)))],
                flags=0x20 | 1, extra=bytes(16))                       # times stored, 2-byte chunk size
    ugroup = ohdr([msg(0x2, struct.pack("<BBQQ", 0, 0, h5.UNDEF, h5.UNDEF)), msg(0x6, link("data", dset))])
    # clock: a scalar whose datatype message is shared — it points at a committed datatype object
    committed = ohdr([msg(0x3, h5._encode_datatype(clock_dtype))])
    shared = struct.pack("<BBQ", 3, 2, committed)
    cdata = alloc(clock.tobytes())
    cset = ohdr([msg(0x1, struct.pack("<BBBB", 2, 0, 0, 0)), msg(0x3, shared, 0x02), msg(0x8, struct.pack("<BBQQ", 3, 1, cdata, clock.nbytes))])
    # the model group: creation order tracked (6-byte message headers), its second link in a continuation chunk
    chunk = alloc(b"OCHK" + msg(0x6, link("clock", cset, wide=True), order=1) + b"\x00" * 4)
    chunk_len = len(buf) - base - chunk
    model = ohdr([msg(0x6, link("u", ugroup), order=0), msg(0x10, struct.pack("<QQ", chunk, chunk_len), order=2), msg(0x0, b"", order=3)],
                 flags=0x04)
    root = ohdr([msg(0x6, link("NonhydrostaticModel", model))])
    sb = h5.SIGNATURE + struct.pack("<BBBBQQQQ", 2, 8, 8, 0, base, h5.UNDEF, len(buf) - base, root) + b"\x00" * 4
    buf[base + superblock_at:base + superblock_at + len(sb)] = sb
    with open(path, "wb") as fh:
        fh.write(buf)


def test_version_2_structures_are_parsed_as_the_specification_lays_them_out(tmp_path):
    rng = np.random.default_rng(5)
    u = np.asfortranarray(rng.random((6, 5, 4)))
    clock = np.array((12.0, 6.0, 6.0, 2, 1), dtype=ck.CLOCK_DTYPE)
    path = tmp_path / "checkpoint_iteration2.jld2"
    _v2_file(path, u, ck.CLOCK_DTYPE, clock)
    f = h5.File(path)
    assert f.superblock_version == 2 and f.base == 512
    assert f.keys() == ["NonhydrostaticModel"] and f["NonhydrostaticModel"].keys() == ["u", "clock"]
    got = f.read("NonhydrostaticModel/u/data")
    assert got.shape == (6, 5, 4) and np.array_equal(got, u)
    assert f.read("NonhydrostaticModel/clock") == {"time": 12.0, "last_Δt": 6.0, "last_stage_Δt": 6.0, "iteration": 2, "stage": 1}


def test_regression_file_layout_round_trip(tmp_path):
    """The layout of the reference's golden regression files (test_nonhydrostatic_regression.jl:7-42) written from a model and read
    back by get_fields_from_checkpoint."""
    view = _ModelView(_thermal_bubble("QuasiAdamsBashforth2"))
    _run(view, 6.0, 3)
    path = tmp_path / "thermal_bubble_regression.h5"
    h5.write(path, ck.regression_tree(view))
    solution, Gn, Gm = ck.get_fields_from_checkpoint(path)
    assert list(solution) == ["u", "v", "w", "S", "T"]                   # tracers in the file's (sorted) key order
    for n, f in zip(view.m.prognostic_names(), view.m.prognostic_fields()):
        assert np.array_equal(solution[n], f.data)
        assert np.array_equal(Gn[n], view.m.Gn[n].data) and np.array_equal(Gm[n], view.m.Gm[n].data)
    assert np.abs(Gm["T"]).max() > 0                                      # AB2 after three steps: G⁻ is populated

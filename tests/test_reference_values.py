"""Value pins: numbers the REFERENCE holds (the printed outputs of its jldoctest blocks, tests/golden/reference_doctest_values.json)
against the oracle (CPU) and against the CUDA library through the C ABI (the host-side coefficient tables on CPU; grid metrics,
set!, fill_halo_regions! on the GPU).  Full-precision prints are compared bit for bit; show(grid) prints (6 significant digits)
after the same formatting."""
import json
import os

import numpy as np
import pytest

import oracle as o
from oracle.advection import stencil_coefficients

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "reference_doctest_values.json")))
FTS = {"float32": np.float32, "float64": np.float64}


def as_ft(strings, FT):
    """The FT number a Julia repr denotes (repr is the shortest string that round-trips, so parsing it is exact)."""
    return np.array([FT(s) for s in strings], dtype=FT)


def sig6(x):
    return float("%.6g" % float(x))


def stencil_in_print_order(coeff):
    # calc_reconstruction_stencil prints coeff[order - idx + 1] for idx = 1..N (reconstruction_coefficients.jl:186-190)
    return tuple(reversed(coeff))


def oracle_coefficients(FT, shift, buffer):
    # uniform_reconstruction_coefficients — reconstruction_coefficients.jl:138-140
    if shift == "symmetric":
        return stencil_coefficients(FT, buffer - 1, 2 * buffer)
    if buffer == 1:
        return (FT(1),)
    return stencil_coefficients(FT, buffer - 2 if shift == "left" else buffer - 1, 2 * buffer - 1)


def stencil_offsets(shift, buffer):
    N = 2 * buffer if shift == "symmetric" else 2 * buffer - 1
    rng = range(2, N + 2) if shift == "right" else range(1, N + 1)
    return [n - buffer - 1 for n in rng]


@pytest.mark.parametrize("case", GOLD["reconstruction_stencils"]["cases"], ids=lambda c: f"{c['FT']}-{c['shift']}-{c['buffer']}")
def test_oracle_reconstruction_coefficients_equal_the_reference_doctest_bits(case):
    FT = FTS[case["FT"]]
    got = np.array(stencil_in_print_order(oracle_coefficients(FT, case["shift"], case["buffer"])), dtype=FT)
    want = as_ft(case["coefficients"], FT)
    assert got.tobytes() == want.tobytes(), (got, want)
    assert stencil_offsets(case["shift"], case["buffer"]) == case["offsets"]


@pytest.mark.parametrize("case", GOLD["reconstruction_stencils"]["cases"], ids=lambda c: f"{c['FT']}-{c['shift']}-{c['buffer']}")
def test_library_reconstruction_tables_equal_the_reference_doctest_bits(case):
    """The tables csrc/grid_host.cpp generates for the kernels (host-only entry point, no device needed)."""
    import oceananigans_jl_b200 as oc
    FT = FTS[case["FT"]]
    got = np.array(stencil_in_print_order(oc.reconstruction_coefficients(FT, case["shift"], case["buffer"])), dtype=FT)
    want = as_ft(case["coefficients"], FT)
    assert got.tobytes() == want.tobytes(), (got, want)


@pytest.mark.parametrize("FT", [np.float32, np.float64])
def test_library_tables_equal_the_oracle_tables_bit_for_bit(FT):
    import oceananigans_jl_b200 as oc
    for buffer in (1, 2):
        assert oc.reconstruction_coefficients(FT, "symmetric", buffer) == oracle_coefficients(FT, "symmetric", buffer)
    for buffer in (1, 2, 3):
        for shift in ("left", "right"):
            assert oc.reconstruction_coefficients(FT, shift, buffer) == oracle_coefficients(FT, shift, buffer)
    for buffer in (2, 3):
        w = o.WENO(FT, 2 * buffer - 1)
        for r in range(buffer):
            assert oc.reconstruction_coefficients(FT, "weno", buffer, r) == tuple(w.coeff[r])


def test_oracle_weno_stencil_offsets_are_the_reference_ones():
    """load_weno_stencil(buffer, :x) — weno_interpolants.jl:376-384: ψ[i + c] for c = -buffer … buffer-1; the oracle's (and
    the kernels') left-biased reconstruction reads exactly that window (offsets -3 … 1 of it for the left bias, 2 … -2
    for the right one)."""
    from oracle.advection import weno_window_offsets
    assert list(weno_window_offsets(3)) == GOLD["weno_stencil_offsets"]["buffer3"]
    assert list(weno_window_offsets(2)) == GOLD["weno_stencil_offsets"]["buffer2"]


def _fields_md_grid(mod):
    g = GOLD["fields_md_grid"]["grid"]
    return mod.RectilinearGrid(FTS[g["FT"]], size=tuple(g["size"]), halo=tuple(g["halo"]), x=tuple(g["x"]), y=tuple(g["y"]),
                               z=g["z_faces"], topology=tuple(g["topology"]))


def test_oracle_grid_metrics_equal_the_reference_doctest_bits():
    G = GOLD["fields_md_grid"]
    g = _fields_md_grid(o)
    cz = g.coords[2]
    assert cz.dc.r(1, 4).tobytes() == as_ft(G["zspacings_center_1_4"], np.float64).tobytes()
    assert cz.df.r(1, 5).tobytes() == as_ft(G["zspacings_face_1_5"], np.float64).tobytes()
    assert cz.C.r(0, 5).tobytes() == as_ft(G["znodes_center_0_5"], np.float64).tobytes()
    assert g.coords[0].dc == np.float64(G["dx"]) and g.coords[1].dc == np.float64(G["dy"])
    assert sig6(cz.dc.r(1, 4).min()) == float(G["min_dz"]) and sig6(cz.dc.r(1, 4).max()) == float(G["max_dz"])


def _summary_grid(mod, case):
    FT = FTS[case.get("FT", "float64")]
    val = lambda v: {"pi": np.pi, "-pi": -np.pi, "2pi": 2 * np.pi, "4pi": 4 * np.pi}.get(v, v) if isinstance(v, str) else v
    kw = {}
    for name in "xyz":
        if name in case:
            kw[name] = tuple(val(v) for v in case[name])
    if "z_faces" in case:
        kw["z"] = case["z_faces"]
    if case.get("z_function") == "tanh_top":
        sigma, Nz, Lz = case["sigma"], case["size"][-1], case["Lz"]
        kw["z"] = lambda k: -Lz * (1 - np.tanh(sigma * (k - 1) / Nz) / np.tanh(sigma))     # rectilinear_grid.jl:209
    return mod.RectilinearGrid(FT, size=tuple(case["size"]), halo=tuple(case["halo"]), topology=tuple(case["topology"]), **kw)


@pytest.mark.parametrize("case", GOLD["grid_summaries"]["cases"], ids=lambda c: c["source"])
def test_oracle_grid_spacings_equal_the_reference_show_output(case):
    g = _summary_grid(o, case)
    for d, name in enumerate("xyz"):
        if "d" + name in case:
            assert sig6(g.coords[d].dc) == float(case["d" + name]), name
    if "min_dz" in case:
        dz = g.coords[2].dc.r(1, g.Nz)
        assert sig6(dz.min()) == float(case["min_dz"]) and sig6(dz.max()) == float(case["max_dz"])


def test_oracle_set_and_periodic_halo_fill_equal_the_reference_doctest():
    from oracle.fields import Field, fill_halo_regions
    H = GOLD["fields_md_halo_fill"]
    g = _fields_md_grid(o)
    assert g.coords[0].C.r(1, 4).tobytes() == as_ft(H["xnodes_center"], np.float64).tobytes()
    c = Field(g, ("c", "c", "c"), name="c")
    c.set(np.broadcast_to((2 * g.coords[0].C.r(1, 4))[:, None, None], c.size))       # set!(c, (x, y, z) -> 2x)
    u = Field(g, ("f", "c", "c"), name="u")
    u.set(np.broadcast_to((2 * g.coords[0].F.r(1, 4))[:, None, None], u.size))
    assert np.array(c.interior[:, 0, 0]).tobytes() == as_ft(H["c_1_4_1_1"], np.float64).tobytes()
    assert np.array(u.interior[:, 0, 0]).tobytes() == as_ft(H["u_1_4_1_1"], np.float64).tobytes()
    before = np.array([[float(v) for v in row] for row in H["before_fill_k1"]])
    after = np.array([[float(v) for v in row] for row in H["after_fill_k1"]])
    k1 = lambda f: np.array(f.at((0, 5), (0, 6), (1, 1)))[:, :, 0]
    assert np.array_equal(k1(c), before)
    fill_halo_regions(c)
    assert np.array_equal(k1(c), after)


# ---- the CUDA library against the same reference-held values (through the C ABI) ------------------------------------------
@pytest.fixture(scope="module")
def oc():
    import oceananigans_jl_b200 as oc
    assert oc.device_count() >= 1
    return oc


@pytest.mark.gpu
def test_library_grid_metrics_equal_the_reference_doctest_bits(oc):
    G = GOLD["fields_md_grid"]
    g = _fields_md_grid(oc)
    dx, dy, dzc, dzf = g.spacings()                          # z arrays start at k = 1 - Hz = 0
    assert dzc[1:5].tobytes() == as_ft(G["zspacings_center_1_4"], np.float64).tobytes()
    assert dzf[1:6].tobytes() == as_ft(G["zspacings_face_1_5"], np.float64).tobytes()
    assert dx == np.float64(G["dx"]) and dy == np.float64(G["dy"])


@pytest.mark.gpu
@pytest.mark.parametrize("case", GOLD["grid_summaries"]["cases"], ids=lambda c: c["source"])
def test_library_grid_spacings_equal_the_reference_show_output(oc, case):
    g = _summary_grid(oc, case)
    dx, dy, dzc, dzf = g.spacings()
    got = {"x": dx, "y": dy, "z": dzc[g.H[2]] if len(dzc) else None}
    for name in "xyz":
        if "d" + name in case:
            assert sig6(got[name]) == float(case["d" + name]), name
    if "min_dz" in case:
        dz = dzc[g.H[2]: g.H[2] + g.N[2]]
        assert sig6(dz.min()) == float(case["min_dz"]) and sig6(dz.max()) == float(case["max_dz"])


@pytest.mark.gpu
def test_library_set_and_periodic_halo_fill_equal_the_reference_doctest(oc):
    H = GOLD["fields_md_halo_fill"]
    g = _fields_md_grid(oc)
    m = oc.NonhydrostaticModel(g, advection=None, tracers=("c",))
    c, u = m.tracers["c"], m.velocities["u"]
    c.set(lambda x, y, z: 2 * x)
    u.set(lambda x, y, z: 2 * x)
    assert c.interior()[:, 0, 0].tobytes() == as_ft(H["c_1_4_1_1"], np.float64).tobytes()
    assert u.interior()[:, 0, 0].tobytes() == as_ft(H["u_1_4_1_1"], np.float64).tobytes()
    before = np.array([[float(v) for v in row] for row in H["before_fill_k1"]])
    after = np.array([[float(v) for v in row] for row in H["after_fill_k1"]])
    assert m.halo == (1, 1, 1)
    assert np.array_equal(c.parent()[:, :, 1], before)       # parent index 1 = k = 1 with Hz = 1
    oc.fill_halo_regions(c)
    assert np.array_equal(c.parent()[:, :, 1], after)


# ---- constants and constructor defaults the reference states ---------------------------------------------------------------
def test_fplane_from_latitude_matches_the_reference_doctest():
    import oceananigans_jl_b200 as oc
    for mod in (o, oc):
        for case in GOLD["coriolis"]["cases"]:
            c = mod.FPlane(case["f"]) if "f" in case else mod.FPlane(rotation_rate=case["rotation_rate"], latitude=case["latitude"])
            assert sig6(c.f) == float(case["printed"]), (mod.__name__, case)
        with pytest.raises(ValueError):
            mod.FPlane()
        with pytest.raises(ValueError):
            mod.FPlane(1e-4, latitude=45)


def test_time_stepper_and_closure_defaults_match_the_reference_constructors():
    from fractions import Fraction
    from oracle.grid import round_fraction
    K = GOLD["constants"]
    g = o.RectilinearGrid(np.float64, size=(4, 4, 4), extent=(1, 1, 1))
    m = o.NonhydrostaticModel(g)
    for key, name in (("gamma1", "g1"), ("gamma2", "g2"), ("gamma3", "g3"), ("zeta2", "z2"), ("zeta3", "z3")):
        for FT in (np.float64, np.float32):
            want = round_fraction(Fraction(*K["rk3"][key]), FT)
            if FT == np.float64:
                assert m.rk3[name] == want, key
    b = o.SeawaterBuoyancy()
    assert (b.g, b.alpha, b.beta) == (float(K["g_Earth"]), float(K["thermal_expansion"]), float(K["haline_contraction"]))
    import oceananigans_jl_b200 as oc
    b1 = oc.SeawaterBuoyancy()
    assert (b1.g, b1.alpha, b1.beta) == (b.g, b.alpha, b.beta)
    c = oc.SmagorinskyLilly()
    assert (c.C, c.Cb, c.Pr) == (float(K["smagorinsky_C"]), float(K["smagorinsky_Cb"]), float(K["smagorinsky_Pr"]))
    assert o.WENO(np.float64, 5).eps == np.float64(np.float32(float(K["weno_epsilon_float32"])))
    assert np.float64(oc.reconstruction_coefficients(np.float64, "weno_optimal", 3)[0]) == np.float64(0.3)

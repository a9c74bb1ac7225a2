# OceananigansB200Ext.jl — the reference-side binding of libnhstep.so (include/nhstep.h) for samlewin/Oceananigans.jl v0.96.19.
#
# The reference attaches backends by dispatch on the architecture type (ext/OceananigansMetalExt.jl:11-35,
# ext/OceananigansReactantExt/TimeSteppers.jl:25,55).  This module does the same for one architecture, `B200`:
#   * `B200NonhydrostaticModel(; grid, advection, buoyancy, coriolis, closure, tracers, timestepper)` builds the Julia
#     `NonhydrostaticModel` on CPU() (clock, closures, boundary-condition objects stay there for callers) AND the `nh_model`
#     behind it, forwarding the constructor arguments of nonhydrostatic_model.jl:114-239 field by field;
#   * `time_step!`, `update_state!`, `compute_tendencies!`, `calculate_pressure_correction!`, `pressure_correct_velocities!`,
#     `fill_halo_regions!`, `set!`, `interior`, `parent`, `solve!` forward to the C ABI (SURVEY.md §8b).
# Julia is not installed in the image this repository is built in, so this file is UNVERIFIED HERE; it mirrors the Python
# binding (`oceananigans.jl_b200/_lib.py`, `model.py`) call for call, that binding is what the GPU tests run, and
# `tests/abi_smoke.c` checks the two struct layouts below from a second language.
module OceananigansB200Ext

using Oceananigans
using Oceananigans.Architectures: AbstractArchitecture
using Oceananigans.Grids: halo_size, topology, znodes, Face
using Oceananigans.Advection: Centered, UpwindBiased, WENO, FluxFormAdvection
using Oceananigans.BuoyancyFormulations: BuoyancyTracer, SeawaterBuoyancy, LinearEquationOfState, BuoyancyForce
using Oceananigans.TurbulenceClosures: ScalarDiffusivity, SmagorinskyLilly, AnisotropicMinimumDissipation,
                                       VerticallyImplicitTimeDiscretization, time_discretization
import Oceananigans.TimeSteppers: time_step!, update_state!
import Oceananigans.Models.NonhydrostaticModels: calculate_pressure_correction!, pressure_correct_velocities!, compute_tendencies!
import Oceananigans.BoundaryConditions: fill_halo_regions!
import Oceananigans.Solvers: solve!
import Oceananigans.Fields: set!, interior

const libnhstep = get(ENV, "NHSTEP_LIB", "libnhstep.so")
const NH_MAX_TRACERS = 8

struct B200 <: AbstractArchitecture; device::Int32; end          # one architecture, no dispatch inside the library
B200() = B200(Int32(0))

# == nh_grid_desc (include/nhstep.h): 4 + 12 + 12 + 12 bytes of Int32 (40), then 6 doubles at offset 40, pointer at 88; 96 bytes
struct NhGridDesc
    float_type::Int32
    N::NTuple{3,Int32}
    H::NTuple{3,Int32}
    topology::NTuple{3,Int32}
    x::NTuple{2,Float64}
    y::NTuple{2,Float64}
    z::NTuple{2,Float64}
    z_faces::Ptr{Float64}
end

# == nh_model_desc (include/nhstep.h), field for field.  Julia lays isbits structs out like C (natural alignment), so the
# paddings fall where the C compiler puts them: 7 Int32 (28 bytes) + 4 padding before ab2_chi, 4 Int32 (16) before the three
# Float64 of the equation of state, has_coriolis + 4 padding, has_closure + 4 padding, vertically_implicit + 4 padding,
# two trailing Int32.  sizeof(NhModelDesc) == 344 == sizeof(nh_model_desc) (tests/abi_smoke.c prints the C side).
struct NhModelDesc
    advection_kind::NTuple{3,Int32}
    advection_order::NTuple{3,Int32}
    timestepper::Int32
    ab2_chi::Float64
    n_tracers::Int32
    buoyancy_kind::Int32
    buoyancy_tracer::NTuple{2,Int32}
    gravitational_acceleration::Float64
    thermal_expansion::Float64
    haline_contraction::Float64
    has_coriolis::Int32
    coriolis_f::Float64
    has_closure::Int32
    nu::Float64
    kappa::NTuple{NH_MAX_TRACERS,Float64}
    vertically_implicit::Int32
    smagorinsky_C::Float64
    smagorinsky_Cb::Float64
    smagorinsky_Pr::NTuple{NH_MAX_TRACERS,Float64}
    amd_Cnu::Float64
    amd_Ckappa::NTuple{NH_MAX_TRACERS,Float64}
    hydrostatic_pressure_anomaly::Int32
    exact_arithmetic::Int32
end

struct NhClock; time::Float64; iteration::Int64; stage::Int32; last_dt::Float64; last_stage_dt::Float64; end

check(status) = status == 0 || error(unsafe_string(ccall((:nh_last_error, libnhstep), Cstring, ())))

topo_code(::Type{Periodic}) = Int32(0); topo_code(::Type{Bounded}) = Int32(1); topo_code(::Type{Flat}) = Int32(2)

# RectilinearGrid(arch, FT; size, x, y, z, halo, topology) — src/Grids/rectilinear_grid.jl:248-275
function nh_grid(grid::RectilinearGrid{FT,TX,TY,TZ}, arch::B200) where {FT,TX,TY,TZ}
    faces = grid.Δzᵃᵃᶜ isa Number ? nothing : Float64.(collect(znodes(grid, Face())))   # stretched z → Nz+1 faces
    h = Ref{Ptr{Cvoid}}()
    GC.@preserve faces begin
        zf = faces === nothing ? Ptr{Float64}(C_NULL) : pointer(faces)
        desc = Ref(NhGridDesc(FT === Float64 ? 1 : 0, Int32.(size(grid)), Int32.(halo_size(grid)),
                              (topo_code(TX), topo_code(TY), topo_code(TZ)),
                              (grid.xᶠᵃᵃ[1], grid.xᶠᵃᵃ[grid.Nx+1]), (grid.yᵃᶠᵃ[1], grid.yᵃᶠᵃ[grid.Ny+1]),
                              (grid.zᵃᵃᶠ[1], grid.zᵃᵃᶠ[grid.Nz+1]), zf))
        check(ccall((:nh_grid_create, libnhstep), Int32, (Ref{NhGridDesc}, Int32, Ref{Ptr{Cvoid}}), desc, arch.device, h))
    end
    return h[]
end

# advection → (kind, order): NH_ADV_NONE 0, CENTERED 1, UPWIND 2, WENO 3
scheme_code(::Nothing) = (Int32(0), Int32(0))
scheme_code(::Centered{B}) where B = (Int32(1), Int32(2B))
scheme_code(::UpwindBiased{B}) where B = (Int32(2), Int32(2B - 1))
scheme_code(::WENO{B}) where B = (Int32(3), Int32(2B - 1))
per_direction(a::FluxFormAdvection) = (a.x, a.y, a.z)
per_direction(a) = (a, a, a)

padded(t, fill) = ntuple(i -> i <= length(t) ? Float64(t[i]) : Float64(fill), NH_MAX_TRACERS)
by_tracer(x::Number, names) = ntuple(_ -> x, length(names))
by_tracer(x::NamedTuple, names) = ntuple(i -> x[names[i]], length(names))

mutable struct B200Model
    handle::Ptr{Cvoid}
    grid_handle::Ptr{Cvoid}
    julia_model            # the CPU() NonhydrostaticModel: clock, closure, boundary-condition objects for callers
    tracer_names::Tuple
end

"""
    B200NonhydrostaticModel(arch = B200(); grid, advection, buoyancy, coriolis, closure, tracers, timestepper, kw...)

Forwards `NonhydrostaticModel(; grid, advection, buoyancy, coriolis, closure, tracers, timestepper)`
(nonhydrostatic_model.jl:114-239) to `nh_model_create`.
"""
function B200NonhydrostaticModel(arch::B200 = B200(); grid, advection = Centered(), buoyancy = nothing, coriolis = nothing,
                                 closure = nothing, tracers = (), timestepper = :RungeKutta3, exact_arithmetic = true, kw...)
    jm = NonhydrostaticModel(; grid, advection, buoyancy, coriolis, closure, tracers, timestepper, kw...)
    names = Tuple(keys(jm.tracers))
    codes = map(scheme_code, per_direction(jm.advection))
    b = buoyancy isa BuoyancyForce ? buoyancy.formulation : buoyancy
    bkind, btr, g, α, β = Int32(0), (Int32(0), Int32(0)), 0.0, 0.0, 0.0
    if b isa BuoyancyTracer
        bkind, btr = Int32(1), (Int32(findfirst(==(:b), names) - 1), Int32(0))
    elseif b isa SeawaterBuoyancy
        b.equation_of_state isa LinearEquationOfState || error("libnhstep: SeawaterBuoyancy needs a LinearEquationOfState")
        bkind = Int32(2)
        btr = (Int32(findfirst(==(:T), names) - 1), Int32(findfirst(==(:S), names) - 1))
        g, α, β = b.gravitational_acceleration, b.equation_of_state.thermal_expansion, b.equation_of_state.haline_contraction
    end
    ckind, ν, κ, vitd = Int32(0), 0.0, padded((), 0), Int32(0)
    C, Cb, Pr, Cν, Cκ = 0.16, 1.0, padded((), 1), 1 / 12, padded((), 1 / 12)
    if closure isa ScalarDiffusivity
        ckind, ν, κ = Int32(1), closure.ν, padded(by_tracer(closure.κ, names), 0)
        vitd = Int32(time_discretization(closure) isa VerticallyImplicitTimeDiscretization)
    elseif closure isa AnisotropicMinimumDissipation
        ckind, Cν, Cκ = Int32(4), closure.Cν, padded(by_tracer(closure.Cκ, names), 1 / 12)
    elseif closure !== nothing                       # Smagorinsky(coefficient = number) / SmagorinskyLilly
        lilly = hasproperty(closure.coefficient, :smagorinsky)
        ckind = lilly ? Int32(3) : Int32(2)
        C = lilly ? closure.coefficient.smagorinsky : closure.coefficient
        Cb = lilly ? closure.coefficient.reduction_factor : 0.0
        Pr = padded(by_tracer(closure.Pr, names), 1)
    end
    desc = Ref(NhModelDesc(map(first, codes), map(last, codes), timestepper == :RungeKutta3 ? 0 : 1,
                           timestepper == :RungeKutta3 ? 0.1 : jm.timestepper.χ, length(names), bkind, btr, g, α, β,
                           coriolis === nothing ? 0 : 1, coriolis === nothing ? 0.0 : coriolis.f, ckind, ν, κ, vitd, C, Cb, Pr, Cν, Cκ,
                           Int32(-1), Int32(exact_arithmetic)))
    gh = nh_grid(grid, arch)
    h = Ref{Ptr{Cvoid}}()
    check(ccall((:nh_model_create, libnhstep), Int32, (Ptr{Cvoid}, Ref{NhModelDesc}, Ref{Ptr{Cvoid}}), gh, desc, h))
    m = B200Model(h[], gh, jm, names)
    push_weno_z_coefficients!(h[], per_direction(jm.advection)[3], eltype(grid))
    finalizer(m) do x
        ccall((:nh_model_destroy, libnhstep), Int32, (Ptr{Cvoid},), x.handle)
        ccall((:nh_grid_destroy, libnhstep), Int32, (Ptr{Cvoid},), x.grid_handle)
    end
    return m
end

# WENO(grid) on a stretched z (weno_reconstruction.jl:98-124): the scheme's and its buffer scheme's per-index tables
# coeff_zᵃᵃᶠ / coeff_zᵃᵃᶜ[r + 2][i], r = −1 … B−1, i = 0 … Nz+1 (reconstruction_coefficients.jl:306-324) → [t, r + 2, i + 1] arrays
push_weno_z_coefficients!(handle, scheme, FT) = nothing
function push_weno_z_coefficients!(handle, scheme::WENO{B}, FT) where B
    for (location, tables) in enumerate((scheme.coeff_zᵃᵃᶠ, scheme.coeff_zᵃᵃᶜ))
        tables === nothing && continue
        n = length(parent(tables[1]))
        a = Array{FT}(undef, B, B + 1, n)
        for r in 1:B+1, i in 1:n, t in 1:B
            a[t, r, i] = parent(tables[r])[i][t]
        end
        check(ccall((:nh_model_set_weno_z_coefficients, libnhstep), Int32, (Ptr{Cvoid}, Int32, Int32, Ptr{Cvoid}, Int64, Int32),
                    handle, B, location - 1, a, length(a), 0))
    end
    push_weno_z_coefficients!(handle, scheme.buffer_scheme, FT)
end

field_id(m::B200Model, name::Symbol) =
    name === :u ? 0 : name === :v ? 1 : name === :w ? 2 : name === :pNHS ? 16 : name === :pHY′ ? 17 :
    3 + findfirst(==(name), m.tracer_names) - 1

# time_step!(model, Δt; euler) — src/TimeSteppers/runge_kutta_3.jl:77, quasi_adams_bashforth_2.jl:74
time_step!(m::B200Model, Δt; callbacks=[], euler=false) =
    check(ccall((:nh_time_step, libnhstep), Int32, (Ptr{Cvoid}, Float64, Int32, Int32), m.handle, Δt, euler, 1))
# update_state!(model; compute_tendencies) — update_nonhydrostatic_model_state.jl:20
update_state!(m::B200Model, callbacks=[]; compute_tendencies=true) =
    check(ccall((:nh_update_state, libnhstep), Int32, (Ptr{Cvoid}, Int32), m.handle, compute_tendencies))
compute_tendencies!(m::B200Model, callbacks) = check(ccall((:nh_compute_tendencies, libnhstep), Int32, (Ptr{Cvoid},), m.handle))
calculate_pressure_correction!(m::B200Model, Δt) =
    check(ccall((:nh_calculate_pressure_correction, libnhstep), Int32, (Ptr{Cvoid}, Float64), m.handle, Δt))
pressure_correct_velocities!(m::B200Model, Δt) =
    check(ccall((:nh_pressure_correct_velocities, libnhstep), Int32, (Ptr{Cvoid}, Float64), m.handle, Δt))
# fill_halo_regions!(fields…) — src/Fields/field_tuples.jl:56; mask bit f = field id f of include/nhstep.h
fill_halo_regions!(m::B200Model, mask::UInt32; fill_boundary_normal_velocities=true) =
    check(ccall((:nh_fill_halo_regions, libnhstep), Int32, (Ptr{Cvoid}, UInt32, Int32), m.handle, mask, fill_boundary_normal_velocities))

# set!(model; u=…, …) — set_nonhydrostatic_model.jl:33-60: interiors go in through nh_set_field, then the projection tail
function set!(m::B200Model; enforce_incompressibility=true, kwargs...)
    FT = eltype(m.julia_model.grid)
    for (name, value) in kwargs
        id = field_id(m, name)
        a = Array{FT}(value)                                       # interior(field) layout, column-major
        check(ccall((:nh_set_field, libnhstep), Int32, (Ptr{Cvoid}, Int32, Int32, Ptr{Cvoid}, Int64, Int32),
                    m.handle, id, 0, a, length(a), 0))
        check(ccall((:nh_fill_halo_regions, libnhstep), Int32, (Ptr{Cvoid}, UInt32, Int32), m.handle, UInt32(1) << id, 1))
    end
    check(ccall((:nh_set_state_finalize, libnhstep), Int32, (Ptr{Cvoid}, Int32), m.handle, enforce_incompressibility))
end

# interior(field) / parent(field): nh_field_size + nh_get_field(region = 0 | 1) into a Julia Array
function field_array(m::B200Model, name::Symbol, region::Integer)
    sz = zeros(Int64, 3)
    check(ccall((:nh_field_size, libnhstep), Int32, (Ptr{Cvoid}, Int32, Int32, Ptr{Int64}), m.handle, field_id(m, name), region, sz))
    a = Array{eltype(m.julia_model.grid)}(undef, sz...)
    check(ccall((:nh_get_field, libnhstep), Int32, (Ptr{Cvoid}, Int32, Int32, Ptr{Cvoid}, Int64, Int32),
                m.handle, field_id(m, name), region, a, length(a), 0))
    return a
end
interior(m::B200Model, name::Symbol) = field_array(m, name, 0)
Base.parent(m::B200Model, name::Symbol) = field_array(m, name, 1)

function clock(m::B200Model)
    c = Ref{NhClock}()
    check(ccall((:nh_clock_get, libnhstep), Int32, (Ptr{Cvoid}, Ref{NhClock}), m.handle, c))
    return c[]
end

# FFTBasedPoissonSolver(grid) / FourierTridiagonalPoissonSolver(grid) and solve!(ϕ, solver, b, m) —
# src/Solvers/fft_based_poisson_solver.jl:52-125, fourier_tridiagonal_poisson_solver.jl:71-147
struct B200PoissonSolver; handle::Ptr{Cvoid}; end
function B200PoissonSolver(grid::RectilinearGrid, arch::B200 = B200(); tridiagonal = !(grid.Δzᵃᵃᶜ isa Number))
    h = Ref{Ptr{Cvoid}}()
    check(ccall((:nh_poisson_create, libnhstep), Int32, (Ptr{Cvoid}, Int32, Ref{Ptr{Cvoid}}), nh_grid(grid, arch), tridiagonal ? 1 : 0, h))
    return B200PoissonSolver(h[])
end
solve!(ϕ::Array, s::B200PoissonSolver, b::Array, m = 0) =
    check(ccall((:nh_poisson_solve, libnhstep), Int32, (Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Float64, Int32), s.handle, b, ϕ, m, 0))

# Checkpoints written by oceananigans.jl_b200/checkpointer.py (classic HDF5 with the reference Checkpointer's addresses,
# src/OutputWriters/checkpointer.jl:223-282): set!(model, filepath) for a Julia NonhydrostaticModel, read with HDF5.jl.
# `data` datasets have the Julia size of parent(field); the clock is stored member by member.
function set_from_b200_checkpoint!(model::NonhydrostaticModel, filepath::AbstractString; h5open)   # pass HDF5.h5open
    h5open(filepath, "r") do file
        g = file["NonhydrostaticModel"]
        for (name, field) in pairs(Oceananigans.prognostic_fields(model))
            haskey(g, string(name)) || (@warn "Field $name does not exist in checkpoint and could not be restored."; continue)
            copyto!(parent(field), read(g["$name/data"]))
            if haskey(g, "timestepper") && haskey(g["timestepper/Gⁿ"], string(name))
                copyto!(parent(model.timestepper.Gⁿ[name]), read(g["timestepper/Gⁿ/$name/data"]))
                copyto!(parent(model.timestepper.G⁻[name]), read(g["timestepper/G⁻/$name/data"]))
            end
        end
        c = read(g["clock"])                                   # compound scalar → NamedTuple (time, last_Δt, last_stage_Δt, iteration, stage)
        model.clock.iteration = c.iteration
        model.clock.time = c.time
        model.clock.last_Δt = c.last_Δt
    end
    return nothing
end

end # module

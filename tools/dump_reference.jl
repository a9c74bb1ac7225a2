# Run with Julia + Oceananigans v0.96.19 on any machine to produce value vectors for the parity configurations
# (SURVEY.md §8c residual risk).  Writes u, v, w, p after 100 steps as raw Float64 column-major files readable by
# numpy.fromfile(..., dtype="<f8").reshape(N, N, N, order="F").  ICs are read from files written by
# tools/make_ics.py so that both sides start from bit-identical arrays.
using Oceananigans
N = parse(Int, get(ENV, "N", "32"))
grid = RectilinearGrid(CPU(), Float64, size=(N, N, N), extent=(2π, 2π, 2π), topology=(Periodic, Periodic, Periodic))
model = NonhydrostaticModel(; grid, advection=WENO(order=5), timestepper=:RungeKutta3)
rd(name) = reshape(reinterpret(Float64, read("ics_$(name)_$(N).bin")), N, N, N)
set!(model, u=rd("u"), v=rd("v"), w=rd("w"))
Δt = 0.2 * (2π / N) / 2.0
for _ in 1:100
    time_step!(model, Δt)
end
for (name, f) in (("u", model.velocities.u), ("v", model.velocities.v), ("w", model.velocities.w), ("p", model.pressures.pNHS))
    write("ref_$(name)_$(N).bin", Array(interior(f)))
end

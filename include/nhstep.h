/* nhstep.h — C ABI of libnhstep.so: a B200 (sm_100a) implementation of the Oceananigans
 * NonhydrostaticModel time step.  Plain C, no torch / C++ types in any signature.
 *
 * Every entry point names the reference interface (samlewin/Oceananigans.jl v0.96.19, paths relative
 * to the reference root) that a Julia `ccall` shim would forward to it; see INTEGRATION.md.
 *
 * Conventions
 *   - every function returns int32_t: 0 = ok, non-zero = error; nh_last_error() gives the message
 *     (thread-local).  No exception crosses the ABI.  There is no CPU fallback: if no CUDA device is
 *     usable, the create calls fail.
 *   - the library owns all device memory.  Host pointers are borrowed for the duration of a call.
 *   - one nh_model = one device + one stream.  Calls on one handle are not thread-safe.  All calls are
 *     asynchronous with respect to the host except nh_get_*, nh_synchronize and calls taking host pointers.
 *   - arrays cross the ABI in the reference's layout: column-major, x fastest, either the field's
 *     `interior` (Nx,Ny,Nz [+1 on a Bounded Face dim]) or its `parent` (interior + 2H per non-Flat dim),
 *     element type = the grid's float type (src/Fields/field.jl:19-33, src/Grids/new_data.jl:15-70).
 */
#ifndef NHSTEP_H
#define NHSTEP_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct nh_grid    nh_grid;
typedef struct nh_model   nh_model;
typedef struct nh_poisson nh_poisson;
typedef struct nh_tridiag nh_tridiag;

enum { NH_F32 = 0, NH_F64 = 1 };
enum { NH_PERIODIC = 0, NH_BOUNDED = 1, NH_FLAT = 2 };
enum { NH_CLOSURE_NONE = 0, NH_CLOSURE_SCALAR = 1, NH_CLOSURE_SMAGORINSKY = 2, NH_CLOSURE_SMAGORINSKY_LILLY = 3, NH_CLOSURE_AMD = 4 };
enum { NH_ADV_NONE = 0, NH_ADV_CENTERED = 1, NH_ADV_UPWIND = 2, NH_ADV_WENO = 3 };
enum { NH_RK3 = 0, NH_AB2 = 1 };
enum { NH_BUOYANCY_NONE = 0, NH_BUOYANCY_TRACER = 1, NH_BUOYANCY_SEAWATER_LINEAR = 2 };
enum { NH_HOST = 0, NH_DEVICE = 1 };
enum { NH_INTERIOR = 0, NH_PARENT = 1 };
enum { NH_SOLVER_FFT = 0, NH_SOLVER_FOURIER_TRIDIAGONAL = 1 };

/* field ids for nh_set_field / nh_get_field / halo masks */
enum {
  NH_FIELD_U = 0, NH_FIELD_V = 1, NH_FIELD_W = 2,
  NH_FIELD_TRACER0 = 3,            /* tracer n = NH_FIELD_TRACER0 + n, n < NH_MAX_TRACERS */
  NH_FIELD_PNHS = 16, NH_FIELD_PHY = 17,
  NH_FIELD_NUE = 18,               /* eddy viscosity νₑ (model.diffusivity_fields.νₑ) of the Smagorinsky / AMD closures */
  NH_FIELD_KAPPAE0 = 19,           /* AMD: eddy diffusivity κₑ of tracer n = NH_FIELD_KAPPAE0 + n (… + NH_MAX_TRACERS - 1 = 26) */
  NH_FIELD_GN0 = 32,               /* Gⁿ of prognostic field f = NH_FIELD_GN0 + f */
  NH_FIELD_GM0 = 48                /* G⁻ of prognostic field f = NH_FIELD_GM0 + f */
};
#define NH_MAX_TRACERS 8

/* RectilinearGrid(arch, FT; size, x, y, z, halo, topology) — src/Grids/rectilinear_grid.jl:248-275.
 * x and y must be regular intervals; z is a regular interval (z_faces == NULL) or Nz+1 increasing faces.
 * Flat dimensions: N = 1 is not required (N is the given extent), H is forced to 0, interval ignored. */
typedef struct {
  int32_t float_type;
  int32_t N[3];
  int32_t H[3];
  int32_t topology[3];
  double  x[2], y[2], z[2];
  const double* z_faces;           /* NULL, or Nz+1 host doubles (converted to FT like the reference does) */
} nh_grid_desc;

/* NonhydrostaticModel(; grid, advection, buoyancy, coriolis, closure, tracers, timestepper,
 *                       hydrostatic_pressure_anomaly) — src/Models/NonhydrostaticModels/nonhydrostatic_model.jl:114-239 */
typedef struct {
  int32_t advection_kind[3];       /* per direction (FluxFormAdvection); same value thrice for a plain scheme */
  int32_t advection_order[3];      /* Centered 2|4, UpwindBiased 1|3|5, WENO 3|5 */
  int32_t timestepper;             /* NH_RK3 | NH_AB2 */
  double  ab2_chi;                 /* QuasiAdamsBashforth2 χ (default 0.1) */
  int32_t n_tracers;
  int32_t buoyancy_kind;
  int32_t buoyancy_tracer[2];      /* tracer index of b (TRACER) or of T and S (SEAWATER_LINEAR) */
  double  gravitational_acceleration, thermal_expansion, haline_contraction;
  int32_t has_coriolis;            /* FPlane */
  double  coriolis_f;
  int32_t has_closure;             /* NH_CLOSURE_*: none, ScalarDiffusivity (explicit, isotropic), Smagorinsky(coefficient =
                                      number), SmagorinskyLilly — src/TurbulenceClosures/turbulence_closure_implementations/
                                      Smagorinskys/smagorinsky.jl:88-137, lilly_coefficient.jl:108-139 — or
                                      AnisotropicMinimumDissipation (Cb = nothing) — …/anisotropic_minimum_dissipation.jl:110-208 */
  double  nu;
  double  kappa[NH_MAX_TRACERS];
  int32_t vertically_implicit;     /* ScalarDiffusivity(VerticallyImplicitTimeDiscretization(), ν, κ): z-diffusion by a
                                      tridiagonal solve after each substep — vertically_implicit_diffusion_solver.jl:57-273,
                                      abstract_scalar_diffusivity_closure.jl:216-252; needs a Bounded z */
  double  smagorinsky_C, smagorinsky_Cb;   /* Cˢ (default 0.16) and the Lilly buoyancy multiplier Cb (default 1) */
  double  smagorinsky_Pr[NH_MAX_TRACERS];  /* turbulent Prandtl number of each tracer: κₑ = ℑ(νₑ) / Pr */
  double  amd_Cnu, amd_Ckappa[NH_MAX_TRACERS];   /* Poincaré constants Cν and Cκ of each tracer (default 1/12) */
  int32_t hydrostatic_pressure_anomaly;   /* -1 default (on iff buoyancy), 0 off, 1 on */
  int32_t exact_arithmetic;        /* 1: kernels compiled without FMA contraction, reference expression order (parity
                                      mode); 0: fused fast kernels where available */
} nh_model_desc;

typedef struct { double time; int64_t iteration; int32_t stage; double last_dt; double last_stage_dt; } nh_clock;

const char* nh_last_error(void);
int32_t nh_version(void);
int32_t nh_device_count(int32_t* count);

/* ---- grid ------------------------------------------------------------------------------------------------ */
int32_t nh_grid_create(const nh_grid_desc* desc, int32_t device, nh_grid** out);
int32_t nh_grid_destroy(nh_grid* grid);
/* with_halo(new_halo, grid) — src/Grids/rectilinear_grid.jl (used by inflate_grid_halo_size, nonhydrostatic_model.jl:243-257) */
int32_t nh_grid_info(const nh_grid* grid, nh_grid_desc* out);          /* desc with the final halo; z_faces = NULL */
/* Δzᵃᵃᶜ[k], Δzᵃᵃᶠ[k] for k = 1-Hz … Nz+Hz as doubles (src/Grids/grid_generation.jl:34-135); spacing[0] = Δx, [1] = Δy */
int32_t nh_grid_spacings(const nh_grid* grid, double* dxdy, double* dzc, double* dzf, int64_t nz_entries);
/* uniform_reconstruction_coefficients(FT, Val(bias), buffer) — src/Advection/reconstruction_coefficients.jl:138-140 — and the
 * WENO sub-stencil coefficients / optimal weights (weno_interpolants.jl:77-93): the tables the kernels use, as generated on the
 * host (exact rational -> FT, last = 1 - sum).  which: 0 symmetric (2*buffer values), 1 left, 2 right (2*buffer-1 values),
 * 3 WENO sub-stencil r (buffer values), 4 WENO optimal weights C* (buffer values).  Values are returned as doubles (exact
 * images of the FT numbers); *count = how many.  Host-only: needs no device. */
int32_t nh_reconstruction_coefficients(int32_t float_type, int32_t which, int32_t buffer, int32_t r, double out[8], int32_t* count);

/* ---- model ----------------------------------------------------------------------------------------------- */
int32_t nh_model_create(nh_grid* grid, const nh_model_desc* desc, nh_model** out);
int32_t nh_model_destroy(nh_model* model);
int32_t nh_model_grid_info(const nh_model* model, nh_grid_desc* out);   /* halo after inflate_grid_halo_size */
/* interior(field) / parent(field) in and out; `where` = NH_HOST or NH_DEVICE pointer; count = #elements in data */
int32_t nh_field_size(const nh_model* model, int32_t field, int32_t region, int64_t size[3]);
int32_t nh_set_field(nh_model* model, int32_t field, int32_t region, const void* data, int64_t count, int32_t where);
int32_t nh_get_field(nh_model* model, int32_t field, int32_t region, void* data, int64_t count, int32_t where);
/* set!(model; …) tail — set_nonhydrostatic_model.jl:44-58: fill halos of the set fields is done by nh_set_field(INTERIOR);
 * this call = update_state!(compute_tendencies=false) [+ pressure solve and correction with Δt = 1 + update_state!] */
int32_t nh_set_state_finalize(nh_model* model, int32_t enforce_incompressibility);

/* FieldBoundaryConditions(west|east|south|north|bottom|top = Value/Gradient/FluxBoundaryCondition(number)) passed as
 * NonhydrostaticModel(; boundary_conditions = (name = …,)) — src/BoundaryConditions/boundary_condition.jl:90-122 (number
 * conditions), field_boundary_conditions.jl:199-201 (converted to eltype(grid)), fill_halo_regions_value_gradient.jl:7-99
 * (Value / Gradient: linear extrapolation into the first halo cell), apply_flux_bcs.jl:50-160 (Flux: Gc[1] += q·A/V,
 * Gc[N] -= q·A/V inside compute_tendencies!).  `field` is a prognostic field id, `dim` a Bounded direction in which the
 * field sits at Centers (a wall-normal velocity keeps the impenetrable condition), `side` 0 = west/south/bottom,
 * 1 = east/north/top.  NH_BC_DEFAULT restores the no-flux default.  Call before nh_set_field / nh_set_state_finalize. */
enum { NH_BC_DEFAULT = 0, NH_BC_FLUX = 1, NH_BC_VALUE = 2, NH_BC_GRADIENT = 3 };
int32_t nh_model_set_boundary_condition(nh_model* model, int32_t field, int32_t dim, int32_t side, int32_t kind, double value);

/* The same with an array condition — Value/Gradient/FluxBoundaryCondition(array), boundary_condition.jl:118
 * (getbc = condition[i, j] with the wall's two transverse grid indices).  `data` holds N_e1 × N_e2 values of the model's float
 * type in the reference's (column-major) order: the lower transverse direction runs fastest, N_e the GRID size of that
 * direction (the fill and flux kernels run over `:yz` / `:xz` / `:xy`, fill_halo_regions.jl:354-363).  The values are copied
 * (a later change of the caller's array needs another call).  `where` = NH_HOST or NH_DEVICE. */
int32_t nh_model_set_boundary_condition_array(nh_model* model, int32_t field, int32_t dim, int32_t side, int32_t kind,
                                              const void* data, int64_t count, int32_t where);

/* WENO(grid) — src/Advection/weno_reconstruction.jl:98-124 with compute_reconstruction_coefficients
 * (reconstruction_coefficients.jl:237-324): on a grid stretched in z a WENO scheme built with the grid reconstructs with per-index
 * sub-stencil coefficients (weno_interpolants.jl:124-127); smoothness indicators, optimal weights and the Centered / UpwindBiased
 * schemes it hands down keep uniform coefficients, as in the reference.  `data` holds (Nz + 2) × (buffer + 1) × buffer values of the
 * model's float type: entry [i][r + 1][t], i = 0 … Nz+1 the reconstruction index, r = −1 … buffer−1 the stencil, t the position in the
 * stencil — `coeff_zᵃᵃᶠ[r + 2][i]` (location 0, Face targets) or `coeff_zᵃᵃᶜ[r + 2][i]` (location 1, Center targets), generated on the
 * host (the reference computes them on the CPU too).  Call once per (buffer, location) the scheme uses: buffer 3 and its near-wall
 * buffer scheme 2 for WENO(order = 5).  Without a call the uniform coefficients are used — `WENO()` without a grid.  The model must
 * not be a fused triply periodic one (a stretched z never is). */
int32_t nh_model_set_weno_z_coefficients(nh_model* model, int32_t buffer, int32_t location, const void* data, int64_t count,
                                         int32_t where);

/* fill_halo_regions!(fields…; fill_boundary_normal_velocities) — src/Fields/field_tuples.jl:56-119,
 * src/BoundaryConditions/fill_halo_regions.jl:50-82.  field_mask bit f = field id f (ids < 32). */
int32_t nh_fill_halo_regions(nh_model* model, uint32_t field_mask, int32_t fill_boundary_normal_velocities);
/* update_state!(model; compute_tendencies) — update_nonhydrostatic_model_state.jl:20-57 */
int32_t nh_update_state(nh_model* model, int32_t compute_tendencies);
/* compute_tendencies!(model, callbacks) — compute_nonhydrostatic_tendencies.jl:17-54 */
int32_t nh_compute_tendencies(nh_model* model);
/* rk3_substep!(model, Δt, γ, ζ) — src/TimeSteppers/runge_kutta_3.jl:160-208 (has_zeta = 0 ⇒ first stage) */
int32_t nh_rk3_substep(nh_model* model, double dt, double gamma, double zeta, int32_t has_zeta);
/* ab2_step!(model, Δt, χ) — src/TimeSteppers/quasi_adams_bashforth_2.jl:128-175 */
int32_t nh_ab2_step(nh_model* model, double dt, double chi);
/* cache_previous_tendencies!(model) — src/TimeSteppers/store_tendencies.jl:12-22 */
int32_t nh_cache_previous_tendencies(nh_model* model);
/* calculate_pressure_correction!(model, Δt), pressure_correct_velocities!(model, Δt) — pressure_correction.jl:8-50 */
int32_t nh_calculate_pressure_correction(nh_model* model, double dt);
int32_t nh_pressure_correct_velocities(nh_model* model, double dt);
/* time_step!(model, Δt; euler) — runge_kutta_3.jl:77-151 / quasi_adams_bashforth_2.jl:74-115; n_steps ≥ 1 repeats it */
int32_t nh_time_step(nh_model* model, double dt, int32_t euler, int32_t n_steps);
/* one time_step! with CUDA events between its phases on the model's stream; phase_ms = {tendencies + substep,
 * pressure solve (halo fill, source term, transforms, real store), pressure correction (+ halo fill), remainder} */
int32_t nh_time_step_profiled(nh_model* model, double dt, int32_t euler, double phase_ms[4]);
int32_t nh_clock_get(const nh_model* model, nh_clock* out);
int32_t nh_clock_set(nh_model* model, const nh_clock* in);
/* max |div U| and other diagnostics used by the reference tests (test/utils_for_runtests.jl:125-151) */
int32_t nh_max_abs_divergence(nh_model* model, double* out);
int32_t nh_synchronize(nh_model* model);
/* raw device pointer + strides of a field (for zero-copy torch views and NCCL halo/transposes) */
int32_t nh_field_device_ptr(nh_model* model, int32_t field, void** ptr, int64_t offset_of_111[1], int64_t strides[3]);
int32_t nh_model_set_stream(nh_model* model, void* cuda_stream);
/* number of kernels launched by this model since creation (bench.py's gpu_launches) */
int32_t nh_launch_count(const nh_model* model, int64_t* count);

/* ---- y-slab multi-GPU time step (one process per GPU) ---------------------------------------------------------
 * Replaces, for the triply periodic regular-grid WENO5 configuration on a Distributed(Partition(1, R, 1)) architecture:
 *   fill_halo_regions! of distributed fields   — src/DistributedComputations/halo_communication.jl:114-366
 *   DistributedFFTBasedPoissonSolver solve!    — src/DistributedComputations/distributed_fft_based_poisson_solver.jl:141-178
 *   transpose_z_to_y! / transpose_y_to_z!      — src/DistributedComputations/distributed_transpose.jl:25-191
 * The library packs, transforms and unpacks on the model's stream; the messages themselves (NCCL send/recv and
 * all-to-all over NVLink, MPI in the reference) are issued by the host on the buffers of nh_dist_buffer().  Per stage:
 *   nh_dist_tendency_substep → exchange(v, 1 row) → nh_dist_poisson(0) → all-to-all A→B → nh_dist_poisson(1)
 *   → all-to-all B→A → nh_dist_poisson(2) → exchange(pNHS, 1 row) → nh_dist_pressure_correct → exchange(u,v,w, H rows)
 * where exchange = nh_dist_halo(pack) → send_lo→previous rank's recv_hi, send_hi→next rank's recv_lo → nh_dist_halo(unpack).
 * The spectral buffers hold R blocks of Cx·NyL·Nz complex values; block d of A goes to rank d (and back for B). */
enum { NH_DIST_SEND_LO = 0, NH_DIST_SEND_HI = 1, NH_DIST_RECV_LO = 2, NH_DIST_RECV_HI = 3, NH_DIST_SPECTRAL_A = 4, NH_DIST_SPECTRAL_B = 5 };
/* the model was created on this rank's LOCAL grid (Nx, Ny/R, Nz; y Periodic); y0, y1, Ny describe the global y extent */
int32_t nh_dist_configure(nh_model* model, int32_t rank, int32_t nranks, double y0, double y1, int32_t Ny_global);
/* mode: 0 = not distributed, 1 = fused path (stepped through the nh_dist_* pieces below), 2 = complete path: any model that
 * is not the fused configuration (tracers, buoyancy, Coriolis, closures, boundary conditions, a Bounded — regular or
 * stretched — z with the distributed Fourier-tridiagonal solve of distributed_fft_tridiagonal_solver.jl:260-291).  Such a
 * model needs nh_dist_ipc_export / nh_dist_ipc_import right after nh_dist_configure and is then driven by the ordinary
 * entry points (nh_set_field, nh_set_state_finalize, nh_time_step, …) called in the same order on every rank: its y halo
 * fills read the neighbours' rows over NVLink and its pressure solve transposes through peer stores, each bracketed by the
 * flag barrier of nh_dist_barrier inside the library. */
int32_t nh_dist_mode(nh_model* model, int32_t* mode);
int32_t nh_dist_buffer(nh_model* model, int32_t which, void** device_ptr, int64_t* bytes);
/* field_mask as in nh_fill_halo_regions (u, v, w, pNHS; at most three); `elements` = values per side just packed/unpacked */
int32_t nh_dist_halo(nh_model* model, uint32_t field_mask, int32_t width, int32_t unpack, int64_t* elements);
/* stage 0…2 of RK3 (runge_kutta_3.jl:77-151) or 0 of AB2; compute_tendencies! fused with the substep; updates the clock */
int32_t nh_dist_tendency_substep(nh_model* model, double dt, int32_t stage, int32_t euler, double* stage_dt);
/* The same sweep in two launches, so that communication overlaps computation like the reference's interior / buffer split
 * (src/Models/interleave_communication_and_computation.jl:9-67, NonhydrostaticModels/compute_nonhydrostatic_buffer_tendencies.jl:
 * 10-83): part 1 = the tile rows whose stencils stay inside this slab's own rows — it needs no neighbour's halo row and is issued
 * BEFORE the barrier that closes the previous stage's pressure correction; part 2 = the tile rows next to the two y faces,
 * issued after that barrier; it also swaps the ping-pong buffers and advances the clock.  part 0 = nh_dist_tendency_substep. */
int32_t nh_dist_tendency_substep_part(nh_model* model, double dt, int32_t stage, int32_t euler, int32_t part, double* stage_dt);
/* phase 0: div(U*)/Δt → FFT x → FFT z into A;  1: FFT y, ÷(λx+λy+λz), IFFT y in B;  2: IFFT z, IFFT x from A → pNHS */
int32_t nh_dist_poisson(nh_model* model, int32_t phase, double stage_dt);
int32_t nh_dist_pressure_correct(nh_model* model, double stage_dt);
/* Peer-to-peer mode (NVLink / NVSwitch peer memory; one process per GPU on ONE node).  After every rank has passed the
 * concatenation of all ranks' export blobs (rank order) to nh_dist_ipc_import, the kernels store their transpose blocks
 * and halo rows straight into the destination rank's buffers: nh_dist_poisson(0) scatters into the peers' y-local
 * buffers, nh_dist_poisson(1) scatters back, nh_dist_poisson(2) also writes the next rank's p[j = 0],
 * nh_dist_tendency_substep the previous rank's v*[j = Ny+1], nh_dist_pressure_correct the neighbours' u, v, w halo rows.
 * The host then issues ONE stream-ordered barrier (e.g. a one-element NCCL all-reduce) after each of those five calls
 * instead of the exchanges / all-to-alls listed above.  Must be called before the first step, at the same point on
 * every rank.  blob = NULL returns the required size in *nbytes. */
int32_t nh_dist_ipc_export(nh_model* model, unsigned char* blob, int64_t capacity, int64_t* nbytes);
int32_t nh_dist_ipc_import(nh_model* model, const unsigned char* all_blobs, int64_t nbytes);
/* Stream-ordered barrier of the peer-to-peer group (replaces the MPI/NCCL synchronisation between the pieces of a stage,
 * cf. sync_device! + MPI.Waitall of halo_communication.jl:91-112): one small kernel publishes an epoch into every peer's flag
 * array over NVLink (release, system scope) and waits for all peers' epochs (acquire).  No reference counterpart.
 * The wait is on the GPU: a rank that waits longer than NHSTEP_BARRIER_TIMEOUT_S seconds (environment, default 600, 0 = for ever)
 * for a peer traps its context, like a collective watchdog.  On a complete-path slab model (nh_dist_mode = 2) every call that
 * fills halos (nh_set_field(INTERIOR) + nh_fill_halo_regions, nh_set_state_finalize, nh_time_step, nh_get_field(PARENT) …)
 * issues such barriers and is therefore COLLECTIVE: every rank must make it, in the same order. */
int32_t nh_dist_barrier(nh_model* model);
/* transpose_z_to_y!(…) (direction 0) / transpose_y_to_z!(…) (direction 1) — src/DistributedComputations/distributed_transpose.jl:
 * 25-191 — as pure data movement between the two spectral buffers of nh_dist_buffer over NVLink peer stores (peer-to-peer mode):
 * block d of this rank's source buffer lands, bit for bit, in block `rank` of rank d's destination buffer.  The time step folds
 * these copies into its FFT passes; callers bracket this entry with nh_dist_barrier.  A -> B -> A is the identity
 * (test/test_distributed_transpose.jl:13-54). */
int32_t nh_dist_transpose(nh_model* model, int32_t direction);

/* ---- standalone solvers ---------------------------------------------------------------------------------- */
/* FFTBasedPoissonSolver(grid) — src/Solvers/fft_based_poisson_solver.jl:52-74;
 * FourierTridiagonalPoissonSolver(grid) — src/Solvers/fourier_tridiagonal_poisson_solver.jl:71-117 */
int32_t nh_poisson_create(nh_grid* grid, int32_t kind, nh_poisson** out);
int32_t nh_poisson_destroy(nh_poisson* solver);
/* solve!(ϕ, solver, b, m) — fft_based_poisson_solver.jl:95-125 / fourier_tridiagonal_poisson_solver.jl:119-147 (+ set_source_term!).
 * rhs and phi are real (Nx,Ny,Nz) arrays of the grid's float type; `where` as above.  m is the Helmholtz shift (FFT solver only). */
int32_t nh_poisson_solve(nh_poisson* solver, const void* rhs, void* phi, double m, int32_t where);
int32_t nh_poisson_synchronize(nh_poisson* solver);
int32_t nh_poisson_launch_count(const nh_poisson* solver, int64_t* count);
/* timing helper: n back-to-back solves on device-resident data, elapsed ms from CUDA events on the solver's stream */
int32_t nh_poisson_time_solves(nh_poisson* solver, const void* rhs_dev, void* phi_dev, int32_t n, float* ms);

/* BatchedTridiagonalSolver(grid; lower_diagonal, diagonal, upper_diagonal, tridiagonal_direction) + solve!(ϕ, solver, rhs)
 * — src/Solvers/batched_tridiagonal_solver.jl:11-79,100-235.  a, c: n-1 host values (1-D), b: n host values (1-D) or
 * Nx·Ny·Nz (3-D) when b_is_3d; direction 0|1|2.  rhs/phi real arrays of the grid float type (host). */
int32_t nh_tridiag_create(nh_grid* grid, const double* a, const double* b, const double* c, int32_t b_is_3d,
                          int32_t direction, nh_tridiag** out);
int32_t nh_tridiag_solve(nh_tridiag* solver, const void* rhs, void* phi);
int32_t nh_tridiag_destroy(nh_tridiag* solver);

/* nh_tridiag_solve_static — the same batched tridiagonal sweep as nh_tridiag_solve for callers that may not allocate device
 * memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO counterpart in the
 * reference.  Every array is the caller's DEVICE memory of the given
 * float type: a, c = N[direction]-1 values, b = N[direction] values (or Nx*Ny*Nz, x fastest, when b_is_3d), rhs and phi =
 * Nx*Ny*Nz values (x fastest).  Where the reference's |beta| guard fails phi keeps what the caller's array held.  The scratch
 * of the forward sweep is a file-scope __device__ array of NH_TRIDIAG_STATIC_MAX_ELEMENTS * 8 bytes (8 MiB, held by every GPU
 * that loads the library until exit): Nx*Ny*Nz beyond that limit is rejected with an error.  The call allocates nothing and
 * is asynchronous on `cuda_stream` (NULL = the default stream).  All calls on a device share the scratch array, so calls that
 * overlap on the GPU race: serialise them.  The first call on a device resolves the array's device address
 * (cudaGetSymbolAddress); make that one outside a graph capture. */
#define NH_TRIDIAG_STATIC_MAX_ELEMENTS (1 << 20)
int32_t nh_tridiag_solve_static(int32_t device, int32_t float_type, const int32_t N[3], int32_t direction,
                                const void* a_dev, const void* b_dev, int32_t b_is_3d, const void* c_dev,
                                const void* rhs_dev, void* phi_dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif

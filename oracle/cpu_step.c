/* oracle/cpu_step.c — TEST INFRASTRUCTURE ONLY (checker and timed CPU baseline; never on the product path).
 *
 * Compiled C / OpenMP restatement of the reference's CPU() code path for the benchmark workload of BASELINE.json
 * configs[1]: NonhydrostaticModel on a regular triply-periodic RectilinearGrid, WENO(order = 5) momentum advection,
 * no tracers / buoyancy / Coriolis / closure, RungeKutta3, FFTBasedPoissonSolver, Float64.  It keeps the reference's
 * loop structure — one pass over the grid per reference kernel, every face flux evaluated at both adjacent cells,
 * a full complex-to-complex 3-D transform, G⁻ ← Gⁿ as a copy — so that its time is a fair stand-in for the Julia
 * CPU() architecture, which cannot run on this box (no Julia).  It is pinned against the NumPy oracle
 * (tests/test_cpu_step.py: tendencies bit-exact, fields after full steps to FFT rounding).
 *
 * Reference functions restated (file:line in samlewin/Oceananigans.jl v0.96.19):
 *   fill_halo_regions! periodic                src/BoundaryConditions/fill_halo_regions_periodic.jl:40-111
 *   compute_Gu!/Gv!/Gw!, u/v/w_velocity_tendency
 *                                              src/Models/NonhydrostaticModels/compute_nonhydrostatic_tendencies.jl:57-139
 *                                              src/Models/NonhydrostaticModels/nonhydrostatic_tendency_kernel_functions.jl:47-200
 *   div_𝐯u/v/w                                 src/Advection/momentum_advection_operators.jl:46-83
 *   advective_momentum_flux_*                  src/Advection/upwind_biased_advective_fluxes.jl:23-93
 *   WENO5 (Z weights, stencil selection)       src/Advection/weno_interpolants.jl:70-348,386-526
 *   rk3_substep_field!                         src/TimeSteppers/runge_kutta_3.jl:77-151,194-208
 *   cache_previous_tendencies!                 src/TimeSteppers/store_tendencies.jl:6-22
 *   _compute_source_term!                      src/Models/NonhydrostaticModels/solve_for_pressure.jl:12-17
 *   solve!(ϕ, ::FFTBasedPoissonSolver)         src/Solvers/fft_based_poisson_solver.jl:95-137
 *   _pressure_correct_velocities!              src/Models/NonhydrostaticModels/pressure_correction.jl:31-37
 *   set!(model; …) projection tail             src/Models/NonhydrostaticModels/set_nonhydrostatic_model.jl:50-56
 * FFT: FFTW.jl is not vendored by the reference; the transform here is a self-contained radix-2 Stockham-free
 * (bit-reversal, decimation-in-time) complex FFT for power-of-two sizes, lines batched over OpenMP threads.
 *
 * Layout: x fastest, halo 3, parent extent (N+6)^3 — the reference's `parent(field)`.
 * Build: oracle/Makefile (gcc -O3 -fopenmp -ffp-contract=off: no FMA contraction, like Julia without @fastmath).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define H 3

typedef struct {
    int N[3], P[3];
    long s[3];          /* element strides of the parent array */
    long np;            /* parent elements */
    double d[3];        /* Δx, Δy, Δz */
    double A[3];        /* Ax = Δy·Δz, Ay = Δx·Δz, Az = Δx·Δy */
    double rV;          /* 1 / ((Δx·Δy)·Δz) */
    double csym[4];     /* Centered(4) symmetric coefficients, reference order */
    double cw[3][3];    /* WENO5 sub-stencil coefficients */
    double cstar[3];    /* optimal weights */
    double sm[3][6];    /* smoothness coefficients */
    double eps;
    double g1, g2, g3, z2, z3;
    double *U[3], *p, *Gn[3], *Gm[3];
    double *lam[3];
    double *cs;         /* complex storage, interleaved, N^3 (no halo), x fastest */
    double *tw[3];      /* twiddles exp(-2πi k/N), k < N/2, interleaved */
    int *brev[3];
    double time;
    long iteration;
} cpu_model;

static inline long idx(const cpu_model *m, int i, int j, int k) {   /* 1-based interior indices */
    return (i + H - 1) + m->s[1] * (long)(j + H - 1) + m->s[2] * (long)(k + H - 1);
}

/* ---------------------------------------------------------------- halos (periodic, x then y then z) */
static void fill_halo(const cpu_model *m, double *f) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    const int Px = m->P[0], Py = m->P[1], Pz = m->P[2];
    const long sy = m->s[1], sz = m->s[2];
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < Pz; ++k)
        for (int j = 0; j < Py; ++j) {
            double *r = f + sy * j + sz * k;
            for (int h = 0; h < H; ++h) {
                r[h] = r[Nx + h];
                r[Nx + H + h] = r[H + h];
            }
        }
#pragma omp parallel for schedule(static)
    for (int k = 0; k < Pz; ++k)
        for (int h = 0; h < H; ++h) {
            double *lo = f + sy * h + sz * k, *hi = f + sy * (Ny + H + h) + sz * k;
            memcpy(lo, f + sy * (Ny + h) + sz * k, sizeof(double) * Px);
            memcpy(hi, f + sy * (H + h) + sz * k, sizeof(double) * Px);
        }
#pragma omp parallel for schedule(static)
    for (int h = 0; h < H; ++h) {
        memcpy(f + sz * h, f + sz * (Nz + h), sizeof(double) * Px * Py);
        memcpy(f + sz * (Nz + H + h), f + sz * (H + h), sizeof(double) * Px * Py);
    }
}

/* ---------------------------------------------------------------- WENO5 */
static inline double weno5(const cpu_model *m, const double *q) {
    /* q[0..4]: the five upwind-ordered values; ψ_r = (q[2-r], q[3-r], q[4-r]) */
    double beta[3], alpha[3], pr[3];
    for (int r = 0; r < 3; ++r) {
        const double a = q[2 - r], b = q[3 - r], c = q[4 - r];
        const double *C = m->sm[r];
        beta[r] = a * (C[0] * a + C[1] * b + C[2] * c) + b * (C[3] * b + C[4] * c) + c * c * C[5];
        const double *w = m->cw[r];
        pr[r] = w[0] * a + w[1] * b + w[2] * c;
    }
    const double tau = fabs(beta[0] - beta[2]);
    for (int r = 0; r < 3; ++r) {
        const double x = tau / (beta[r] + m->eps);
        alpha[r] = m->cstar[r] * (1.0 + x * x);
    }
    const double asum = alpha[0] + alpha[1] + alpha[2];
    return alpha[0] / asum * pr[0] + alpha[1] / asum * pr[1] + alpha[2] / asum * pr[2];
}

/* biased interpolation of f along stride s at offset o; s0 = 1 for a Center target, 0 for a Face target */
static inline double biased(const cpu_model *m, const double *f, long o, long s, int s0, int left) {
    double q[5];
    const double *c = f + o + (long)(s0 - 3) * s;      /* S[t] = c[t·s], t = 0..5 */
    if (left) { for (int t = 0; t < 5; ++t) q[t] = c[t * s]; }
    else      { for (int t = 0; t < 5; ++t) q[t] = c[(5 - t) * s]; }
    return weno5(m, q);
}

/* advective_momentum_flux_{U_b}{u_a} at parent offset o (flux location: Center if a == b, Face-Face otherwise) */
static inline double momentum_flux(const cpu_model *m, int a, int b, long o) {
    const double *Ub = m->U[b], *Ua = m->U[a];
    const double A = m->A[b];
    const double *c = m->csym;
    double ut;
    if (a == b) {
        const long s = m->s[b];
        ut = c[3] * (A * Ub[o - s]) + c[2] * (A * Ub[o]) + c[1] * (A * Ub[o + s]) + c[0] * (A * Ub[o + 2 * s]);
        return ut * biased(m, Ua, o, s, 1, ut > 0);
    }
    const long sa = m->s[a];
    ut = c[3] * (A * Ub[o - 2 * sa]) + c[2] * (A * Ub[o - sa]) + c[1] * (A * Ub[o]) + c[0] * (A * Ub[o + sa]);
    return ut * biased(m, Ua, o, m->s[b], 0, ut > 0);
}

/* compute_Gu!/Gv!/Gw!: one kernel per component, each face flux evaluated by both adjacent cells */
static void compute_G(const cpu_model *m, int a) {
    double *G = m->Gn[a];
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j) {
            long o = idx(m, 1, j, k);
            for (int i = 1; i <= Nx; ++i, ++o) {
                double t[3];
                for (int b = 0; b < 3; ++b) {
                    const long s = m->s[b];
                    t[b] = (a == b) ? momentum_flux(m, a, b, o) - momentum_flux(m, a, b, o - s)
                                    : momentum_flux(m, a, b, o + s) - momentum_flux(m, a, b, o);
                }
                G[o] = -(m->rV * (t[0] + t[1] + t[2]));
            }
        }
}

static void compute_tendencies(const cpu_model *m) {
    for (int a = 0; a < 3; ++a) compute_G(m, a);
}

/* ---------------------------------------------------------------- substep / cache */
static void rk3_substep(const cpu_model *m, double dt, double gamma, double zeta, int first) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    for (int a = 0; a < 3; ++a) {
        double *U = m->U[a];
        const double *Gn = m->Gn[a], *Gm = m->Gm[a];
#pragma omp parallel for collapse(2) schedule(static)
        for (int k = 1; k <= Nz; ++k)
            for (int j = 1; j <= Ny; ++j) {
                long o = idx(m, 1, j, k);
                if (first) for (int i = 0; i < Nx; ++i) U[o + i] += dt * gamma * Gn[o + i];
                else       for (int i = 0; i < Nx; ++i) U[o + i] += dt * (gamma * Gn[o + i] + zeta * Gm[o + i]);
            }
    }
}

static void cache_previous_tendencies(const cpu_model *m) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    for (int a = 0; a < 3; ++a) {
#pragma omp parallel for collapse(2) schedule(static)
        for (int k = 1; k <= Nz; ++k)
            for (int j = 1; j <= Ny; ++j) {
                long o = idx(m, 1, j, k);
                memcpy(m->Gm[a] + o, m->Gn[a] + o, sizeof(double) * Nx);
            }
    }
}

/* ---------------------------------------------------------------- FFT (power-of-two, in place, one line) */
static void fft_line(double *x, int n, const double *tw, const int *brev, int inverse) {
    for (int i = 0; i < n; ++i) {
        int j = brev[i];
        if (j > i) {
            double tr = x[2 * i], ti = x[2 * i + 1];
            x[2 * i] = x[2 * j]; x[2 * i + 1] = x[2 * j + 1];
            x[2 * j] = tr; x[2 * j + 1] = ti;
        }
    }
    for (int len = 2; len <= n; len <<= 1) {
        const int half = len >> 1, step = n / len;
        for (int base = 0; base < n; base += len)
            for (int q = 0; q < half; ++q) {
                const double wr = tw[2 * q * step], wi = inverse ? -tw[2 * q * step + 1] : tw[2 * q * step + 1];
                double *a = x + 2 * (base + q), *b = x + 2 * (base + q + half);
                const double br = b[0] * wr - b[1] * wi, bi = b[0] * wi + b[1] * wr;
                b[0] = a[0] - br; b[1] = a[1] - bi;
                a[0] += br; a[1] += bi;
            }
    }
}

/* transform along dimension d of the N^3 complex array (x fastest) */
static void fft_dim(const cpu_model *m, int d, int inverse) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    const int n = m->N[d];
    double *cs = m->cs;
    if (d == 0) {
#pragma omp parallel for schedule(static)
        for (long l = 0; l < (long)Ny * Nz; ++l) fft_line(cs + 2 * l * Nx, Nx, m->tw[0], m->brev[0], inverse);
        return;
    }
    const long stride = d == 1 ? Nx : (long)Nx * Ny;
    const int outer = d == 1 ? Nz : Ny;              /* the other non-x dimension */
    const long ostride = d == 1 ? (long)Nx * Ny : Nx;
    enum { B = 8 };
#pragma omp parallel
    {
        double *buf = (double *)malloc(sizeof(double) * 2 * n * B);
#pragma omp for collapse(2) schedule(static)
        for (int o = 0; o < outer; ++o)
            for (int i0 = 0; i0 < Nx; i0 += B) {
                const int nb = Nx - i0 < B ? Nx - i0 : B;
                double *base = cs + 2 * (o * ostride + i0);
                for (int q = 0; q < n; ++q)
                    for (int c = 0; c < nb; ++c) {
                        buf[2 * (c * n + q)] = base[2 * (q * stride + c)];
                        buf[2 * (c * n + q) + 1] = base[2 * (q * stride + c) + 1];
                    }
                for (int c = 0; c < nb; ++c) fft_line(buf + 2 * c * n, n, m->tw[d], m->brev[d], inverse);
                for (int q = 0; q < n; ++q)
                    for (int c = 0; c < nb; ++c) {
                        base[2 * (q * stride + c)] = buf[2 * (c * n + q)];
                        base[2 * (q * stride + c) + 1] = buf[2 * (c * n + q) + 1];
                    }
            }
        free(buf);
    }
}

/* ---------------------------------------------------------------- pressure */
static void compute_source_term(const cpu_model *m, double dt) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    const double *u = m->U[0], *v = m->U[1], *w = m->U[2];
    const double Ax = m->A[0], Ay = m->A[1], Az = m->A[2];
    const long sy = m->s[1], sz = m->s[2];
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j) {
            long o = idx(m, 1, j, k);
            double *c = m->cs + 2 * ((long)Nx * ((j - 1) + (long)Ny * (k - 1)));
            for (int i = 0; i < Nx; ++i, ++o) {
                const double div = m->rV * ((Ax * u[o + 1] - Ax * u[o]) + (Ay * v[o + sy] - Ay * v[o]) + (Az * w[o + sz] - Az * w[o]));
                c[2 * i] = div / dt;
                c[2 * i + 1] = 0.0;
            }
        }
}

static void poisson_solve(const cpu_model *m) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    for (int d = 0; d < 3; ++d) fft_dim(m, d, 0);
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 0; k < Nz; ++k)
        for (int j = 0; j < Ny; ++j) {
            double *c = m->cs + 2 * ((long)Nx * (j + (long)Ny * k));
            for (int i = 0; i < Nx; ++i) {
                const double den = m->lam[0][i] + m->lam[1][j] + m->lam[2][k];
                c[2 * i] = -c[2 * i] / den;
                c[2 * i + 1] = -c[2 * i + 1] / den;
            }
        }
    m->cs[0] = 0.0; m->cs[1] = 0.0;
    for (int d = 0; d < 3; ++d) fft_dim(m, d, 1);
    const double norm = 1.0 / ((double)Nx * Ny * Nz);
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j) {
            long o = idx(m, 1, j, k);
            const double *c = m->cs + 2 * ((long)Nx * ((j - 1) + (long)Ny * (k - 1)));
            for (int i = 0; i < Nx; ++i) m->p[o + i] = c[2 * i] * norm;
        }
}

static void calculate_pressure_correction(const cpu_model *m, double dt) {
    for (int a = 0; a < 3; ++a) fill_halo(m, m->U[a]);
    compute_source_term(m, dt);
    poisson_solve(m);
    fill_halo(m, m->p);
}

static void pressure_correct_velocities(const cpu_model *m, double dt) {
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    const double *p = m->p;
#pragma omp parallel for collapse(2) schedule(static)
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j) {
            long o = idx(m, 1, j, k);
            for (int i = 0; i < Nx; ++i, ++o)
                for (int a = 0; a < 3; ++a)
                    m->U[a][o] -= (p[o] - p[o - m->s[a]]) / m->d[a] * dt;
        }
}

static void update_state(const cpu_model *m, int tendencies) {
    for (int a = 0; a < 3; ++a) fill_halo(m, m->U[a]);
    if (tendencies) compute_tendencies(m);
}

/* ---------------------------------------------------------------- exported */
cpu_model *cpu_model_create(const int *N, const double *d, const double *lamx, const double *lamy, const double *lamz,
                            const double *csym, const double *cw, const double *cstar, const double *sm, double eps,
                            const double *rk3) {
    for (int a = 0; a < 3; ++a) if (N[a] < 2 || (N[a] & (N[a] - 1))) return NULL;   /* power-of-two sizes only */
    cpu_model *m = (cpu_model *)calloc(1, sizeof(cpu_model));
    const double *lam[3] = {lamx, lamy, lamz};
    for (int a = 0; a < 3; ++a) { m->N[a] = N[a]; m->P[a] = N[a] + 2 * H; m->d[a] = d[a]; }
    m->s[0] = 1; m->s[1] = m->P[0]; m->s[2] = (long)m->P[0] * m->P[1];
    m->np = m->s[2] * m->P[2];
    m->A[0] = d[1] * d[2]; m->A[1] = d[0] * d[2]; m->A[2] = d[0] * d[1];
    m->rV = 1.0 / (m->A[2] * d[2]);
    memcpy(m->csym, csym, sizeof m->csym); memcpy(m->cw, cw, sizeof m->cw);
    memcpy(m->cstar, cstar, sizeof m->cstar); memcpy(m->sm, sm, sizeof m->sm);
    m->eps = eps;
    m->g1 = rk3[0]; m->g2 = rk3[1]; m->g3 = rk3[2]; m->z2 = rk3[3]; m->z3 = rk3[4];
    for (int a = 0; a < 3; ++a) {
        m->U[a] = (double *)calloc(m->np, sizeof(double));
        m->Gn[a] = (double *)calloc(m->np, sizeof(double));
        m->Gm[a] = (double *)calloc(m->np, sizeof(double));
        m->lam[a] = (double *)malloc(sizeof(double) * N[a]);
        memcpy(m->lam[a], lam[a], sizeof(double) * N[a]);
        const int n = N[a];
        m->tw[a] = (double *)malloc(sizeof(double) * n);
        m->brev[a] = (int *)malloc(sizeof(int) * n);
        for (int k = 0; k < n / 2; ++k) {
            const long double ang = -2.0L * 3.14159265358979323846264338327950288L * k / n;
            m->tw[a][2 * k] = (double)cosl(ang); m->tw[a][2 * k + 1] = (double)sinl(ang);
        }
        int bits = 0; while ((1 << bits) < n) ++bits;
        for (int i = 0; i < n; ++i) { int r = 0; for (int b = 0; b < bits; ++b) if (i & (1 << b)) r |= 1 << (bits - 1 - b); m->brev[a][i] = r; }
    }
    m->p = (double *)calloc(m->np, sizeof(double));
    m->cs = (double *)calloc(2 * (size_t)N[0] * N[1] * N[2], sizeof(double));
    return m;
}

void cpu_model_destroy(cpu_model *m) {
    if (!m) return;
    for (int a = 0; a < 3; ++a) { free(m->U[a]); free(m->Gn[a]); free(m->Gm[a]); free(m->lam[a]); free(m->tw[a]); free(m->brev[a]); }
    free(m->p); free(m->cs); free(m);
}

static double *field_ptr(cpu_model *m, int which) {
    if (which < 3) return m->U[which];
    if (which == 3) return m->p;
    if (which < 7) return m->Gn[which - 4];
    return m->Gm[which - 7];
}

/* interior arrays, x fastest (Nx·Ny·Nz doubles) */
void cpu_model_get(cpu_model *m, int which, double *out) {
    const double *f = field_ptr(m, which);
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j)
            memcpy(out + (long)Nx * ((j - 1) + (long)Ny * (k - 1)), f + idx(m, 1, j, k), sizeof(double) * Nx);
}

static void put_interior(cpu_model *m, int which, const double *in) {
    double *f = field_ptr(m, which);
    const int Nx = m->N[0], Ny = m->N[1], Nz = m->N[2];
    for (int k = 1; k <= Nz; ++k)
        for (int j = 1; j <= Ny; ++j)
            memcpy(f + idx(m, 1, j, k), in + (long)Nx * ((j - 1) + (long)Ny * (k - 1)), sizeof(double) * Nx);
}

/* set!(model; u, v, w) incl. the incompressibility projection with Δt = 1 */
void cpu_model_set(cpu_model *m, const double *u, const double *v, const double *w, int enforce_incompressibility) {
    const double *in[3] = {u, v, w};
    for (int a = 0; a < 3; ++a) { put_interior(m, a, in[a]); fill_halo(m, m->U[a]); }
    update_state(m, 0);
    if (enforce_incompressibility) {
        calculate_pressure_correction(m, 1.0);
        pressure_correct_velocities(m, 1.0);
        update_state(m, 0);
    }
}

void cpu_model_compute_tendencies(cpu_model *m) { update_state(m, 1); }

/* time_step!(model, Δt) — RungeKutta3 */
void cpu_model_time_step(cpu_model *m, double dt) {
    if (m->iteration == 0) update_state(m, 1);
    const double gam[3] = {m->g1, m->g2, m->g3}, zet[3] = {0.0, m->z2, m->z3};
    const double sdt[3] = {m->g1 * dt, (m->g2 + m->z2) * dt, (m->g3 + m->z3) * dt};
    for (int st = 0; st < 3; ++st) {
        rk3_substep(m, dt, gam[st], zet[st], st == 0);
        calculate_pressure_correction(m, sdt[st]);
        pressure_correct_velocities(m, sdt[st]);
        if (st < 2) cache_previous_tendencies(m);
        update_state(m, 1);
        m->time += sdt[st];
    }
    m->iteration += 1;
}

/* torchrun exports OMP_NUM_THREADS=1 to its ranks; the baseline is meant to use every host core it may run on */
void cpu_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int cpu_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

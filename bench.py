#!/usr/bin/env python
"""bench.py — cell-updates/s of the NonhydrostaticModel WENO5 RK3 time step (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--size 256] [--impl reference]

A "step" is one time_step!(model, Δt) (3 RK3 substeps incl. 3 pressure solves) on a triply periodic Float64 grid;
one cell-update = one grid cell advanced by one step.  Prints ONE JSON line (see the task contract): `value` is the
device-resident throughput (CUDA events), `e2e` the same step driven through the C ABI with HOST buffers (u, v, w copied
host→device before and u, v, w, p device→host after every step), `roofline` the dominant phase against measured HBM
bandwidth, `cpu_baseline` the compiled C / OpenMP restatement of the reference's CPU path (oracle/cpu_step.c) on a bounded sample,
all host threads.

Beside the contract's keys the N = 1 line carries `targets` (the other BASELINE.json configurations timed in the same process:
the north_star 512³ step, the standalone solvers of configs[2] / configs[3], a 256³ LES-type model) and the N > 1 lines carry
`parity` (10 steps of a small slab problem on the N ranks against the same problem on rank 0 alone) and `targets` (512³ per
GPU, the size SURVEY.md §8e defines the scaling target on, and configs[4]: 1024³ global).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cell-updates/s (NonhydrostaticModel WENO step)"
UNIT = "cell-updates/s"
# FP64-pipe instructions per cell of the fused sweep (ncu opcode mix, profiles/r2_tendency_sass_mix.md: DMUL + DFMA + DADD of
# tendency_fast_kernel<double, 0> in exact and of tendency_plane_kernel<double, 1> in contracted arithmetic)
FP64_INSTR_PER_CELL = {"exact": 1290.0, "contracted": 717.0}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def measured_traffic(phase, N):
    """DRAM bytes per launch of the phase's kernel(s) from the committed ncu capture, if it was taken at this size."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_traffic.json")) as f:
            t = json.load(f)
        return t["dram_bytes_per_launch"][phase] if t["size"] == N else None
    except Exception:
        return None


def smooth_ics(N, seed, size=None, rows=None):
    """Σ of random-phase low-wavenumber sinusoids + 1e-3 noise, built per axis to stay cheap.
    ``size`` = global (Nx, Ny, Nz) (default N³); ``rows`` = (first row, count) restricts y to one rank's slab."""
    rng = np.random.default_rng(seed)
    Nx, Ny, Nz = size or (N, N, N)
    j0, nj = rows or (0, Ny)
    x, y, z = np.arange(Nx) / Nx, (j0 + np.arange(nj)) / Ny, np.arange(Nz) / Nz
    out = {}
    for name in "uvw":
        v = np.zeros((Nx, nj, Nz))
        for _ in range(4):
            k = rng.integers(0, 4, 3)
            ph = rng.uniform(0, 2 * np.pi, 3)
            v += (np.cos(2 * np.pi * k[0] * x + ph[0])[:, None, None] * np.cos(2 * np.pi * k[1] * y + ph[1])[None, :, None]
                  * np.cos(2 * np.pi * k[2] * z + ph[2])[None, None, :])
        v += 1e-3 * np.random.default_rng(seed + 1000 + j0).uniform(-1, 1, (Nx, nj, Nz))
        out[name] = v
    return out


def cpu_oracle_rate(N, steps, warmup=1, budget_s=None):
    """The compiled C / OpenMP restatement of the reference's CPU() path (oracle/cpu_step.c: one pass per reference
    kernel, every face flux evaluated twice, full complex radix-2 FFT of its own — not FFTW — pinned against the NumPy oracle
    in tests/test_cpu_step.py) on an N³ sample of the same workload, on all host threads OpenMP gives it.  With ``budget_s`` the
    timed loop stops early (after at least 2 steps) once that much time has gone.  Returns (rate, s/step, threads, steps)."""
    from oracle.cpu_step import CpuStepModel
    CpuStepModel.use_all_cores()
    m = CpuStepModel((N, N, N), (2 * np.pi,) * 3)
    ics = smooth_ics(N, 1234)
    m.set(ics["u"], ics["v"], ics["w"])
    dt = 0.2 * (2 * np.pi / N) / 2.0
    m.time_step(dt, max(1, warmup))
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        m.time_step(dt, 1)
        done += 1
        if budget_s is not None and done >= 2 and time.perf_counter() - t0 > budget_s:
            break
    t = time.perf_counter() - t0
    return N ** 3 * done / t, t / done, m.threads(), done


CPU_SAMPLE = ("oracle/cpu_step.c (C / OpenMP restatement of the reference CPU() path with its own radix-2 FFT, not the reference's "
              "FFTW; Julia is absent on this box), {n}^3 cells x {k} steps on {c} threads, {s:.3f} s/step")


class NvmlSampler:
    """SM clock and throttle reasons of one GPU every 50 ms from a thread of this process (NVML, the library behind
    nvidia-smi).  A separate `nvidia-smi -lms` process next to the ranks slowed the barrier-synchronised multi-GPU step
    by 20–30 % regardless of its sampling interval, so the in-process reader is preferred."""

    def __init__(self, index):
        import threading
        import pynvml
        self.nv = pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            index = int(vis.split(",")[index])
        self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
        self.samples, self.first, self.run = [], 0, True
        self.t = threading.Thread(target=self._loop, daemon=True)
        self.t.start()

    def _loop(self):
        nv = self.nv
        while self.run:
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:
                pass
            time.sleep(float(os.environ.get("NHSTEP_CLOCK_MS", "50")) * 1e-3)

    def wait_first_sample(self, timeout=5.0):
        t0 = time.time()
        while not self.samples and time.time() - t0 < timeout:
            time.sleep(0.01)

    def mark(self):
        self.first = len(self.samples)

    def stop(self):
        time.sleep(0.06)
        self.run = False
        self.t.join(timeout=2)
        nv, use = self.nv, self.samples[self.first:] or self.samples
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksEventReasonHwThermalSlowdown,
                "sw_thermal_slowdown": nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksEventReasonSwPowerCap}
        reasons = sorted(n for n, b in bits.items() if any(s[2] & b for s in use))
        return {"sm_mhz": float(np.median([s[0] for s in use])), "sm_max_mhz": float(max(s[1] for s in use)),
                "reasons": reasons, "samples": len(use), "source": "nvml"}


def make_sampler(index):
    if os.environ.get("NHSTEP_NO_CLOCKS"):
        return None
    if not os.environ.get("NHSTEP_CLOCK_SMI"):
        try:
            return NvmlSampler(index)
        except Exception:
            pass
    return ClockSampler(index)


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", os.environ.get("NHSTEP_CLOCK_MS", "50")], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def lines(self):
        try:
            with open(self.f.name) as f:
                return sum(1 for _ in f)
        except OSError:
            return 0

    def wait_first_sample(self, timeout=5.0):
        t0 = time.time()
        while self.p is not None and self.lines() == 0 and time.time() - t0 < timeout:
            time.sleep(0.05)

    def mark(self):
        """Call at the start of the timed region: only samples taken after this point are reported."""
        self.first = self.lines()

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.12)          # let the sample covering the end of the timed region land
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        all_lines = self.f.read().splitlines()
        first = getattr(self, "first", 0)
        use = all_lines[first:] if len(all_lines) > first else all_lines
        for line in use:
            parts = [s.strip() for s in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.f.name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(np.max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def workload_config(args, world):
    """The `config` object of the JSON line: a static description of the workload, identical on both arms."""
    N = args.size
    fx, fy = WORLD_FACTORS.get(world, (1, world))
    gsize = (N * fx, N * fy, N)
    exact = args.arithmetic == "exact"
    return {"workload": f"3D triply-periodic NonhydrostaticModel {N}^3 Float64 WENO5 RK3 FFTBasedPoissonSolver"
                        + (f" per GPU (global {gsize[0]}x{gsize[1]}x{gsize[2]})" if world > 1 else ""),
            "arithmetic": ("exact: bit-identical to the reference's CPU arithmetic (no FMA contraction)" if exact else
                           "contracted: FMA contraction + regrouped WENO5 weights in the tendency sweep; rel. L2 vs the compiled "
                           "CPU oracle after 100 steps at this size: profiles/r2_parity_errors.jsonl"),
            "l2_policy": "working set (>= 10 fields x %.0f MB) exceeds the 126 MB L2" % (N ** 3 * 8 / 1e6),
            "parallelism": "single GPU" if world == 1 else
            (f"y-slabs over {world} GPUs: kernels store halo rows and FFT transpose blocks into the peers' buffers over NVLink "
             "(CUDA IPC peer pointers), 5 stream-ordered flag barriers per stage (nh_dist_barrier; no NCCL on the data path)"
             if not os.environ.get("NHSTEP_NO_P2P") else
             f"y-slabs over {world} GPUs: NCCL halo exchange + 2 all-to-all transposes per pressure solve")}


WORLD_FACTORS = {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4), 16: (4, 4)}


def run_reference(args, rank, world):
    """The reference arm: the CPU implementation of the same step on the box's host cores (oracle/cpu_step.c — Julia is
    absent, so `kind` is "port"), rank 0 only, each step one N³ sample of the workload (N = --ref-size, default --size)."""
    if rank != 0:
        return
    N = args.ref_size or args.size
    rate, sps, cores, done = cpu_oracle_rate(N, max(1, args.steps), max(1, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": workload_config(args, world),
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": CPU_SAMPLE.format(n=N, k=done, c=cores, s=sps)},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


FT_B = 8
# algorithmic bytes per cell and step (w = 8 B, F = 3): tendencies + substep 9w + 2·12w, three pressure solves
# 3·12w, three corrections 3·7w = 90w = 720 B
STEP_WORDS = 90


def event_time(fn, stream_sync):
    """ms of fn() between two CUDA events on torch's current stream (the library's default stream is the same stream 0)."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream_sync()
    e0.record()
    fn()
    e1.record()
    stream_sync()
    return e0.elapsed_time(e1)


def solver_target(oc, FT, size, z, reps=10):
    """Standalone solve on device-resident data (BASELINE.json configs[2] / configs[3]): ms per solve from CUDA events on the
    solver's stream (nh_poisson_time_solves) and the HBM fraction for SURVEY.md §8d's 10 words per cell."""
    import ctypes as C
    import torch
    from oceananigans_jl_b200._lib import lib, check
    topo = (oc.Periodic, oc.Periodic, oc.Bounded)
    kw = dict(x=(0, 1), y=(0, 1), z=z)
    g = oc.RectilinearGrid(FT, size=size, topology=topo, **kw)
    solver = (oc.FourierTridiagonalPoissonSolver if g.z_stretched else oc.FFTBasedPoissonSolver)(g)
    tdt = torch.float64 if FT == np.float64 else torch.float32
    rhs = torch.randn(size[::-1], dtype=tdt, device="cuda")
    rhs -= rhs.mean()
    phi = torch.empty_like(rhs)
    ms = C.c_float()
    for n in (3, reps):
        check(lib().nh_poisson_time_solves(solver._h, C.c_void_p(rhs.data_ptr()), C.c_void_p(phi.data_ptr()), n, C.byref(ms)))
    per = ms.value / reps
    cells = size[0] * size[1] * size[2]
    peak, _ = peaks()
    gbs = 10 * np.dtype(FT).itemsize * cells / (per * 1e-3) / 1e9
    return {"ms_per_solve": per, "algorithmic_words_per_cell": 10, "achieved_gbs": gbs, "frac": gbs / peak, "solves": reps}


def tanh_faces(Nz):
    k = np.arange(1, Nz + 2)
    return np.tanh(1.3 * (2 * (k - 1) / Nz - 1)) / np.tanh(1.3)        # test_time_stepping.jl:439-445


def single_gpu_targets(oc, args, exact):
    """The other BASELINE.json configurations, timed in this process on device-resident data (a few seconds each)."""
    import torch
    peak, _ = peaks()
    out = {}

    def guarded(name, fn):
        try:
            out[name] = fn()
        except Exception as e:                                    # a target that cannot run must not cost the headline line
            out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.synchronize()
        torch.cuda.empty_cache()

    def step_512():
        N = 512
        g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(oc.Periodic,) * 3)
        dt = 0.2 * (2 * np.pi / N) / 2.0
        res = {}
        for name, ex in (("contracted", False), ("exact", True)):
            m = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), timestepper="RungeKutta3", exact_arithmetic=ex)
            oc.set_model(m, **smooth_ics(N, 1234))
            oc.time_step(m, dt, n_steps=3)
            n = 10
            ms = event_time(lambda: oc.time_step(m, dt, n_steps=n), torch.cuda.synchronize) / n
            ph = {}
            for _ in range(3):
                for k, v in oc.time_step_profiled(m, dt).items():
                    ph[k] = ph.get(k, 0.0) + v / 3
            gbs = STEP_WORDS * FT_B * N ** 3 / (ms * 1e-3) / 1e9
            res[name] = {"ms_per_step": ms, "value": N ** 3 / (ms * 1e-3), "unit": UNIT, "steps": n, "whole_step_gbs": gbs,
                         "whole_step_frac": gbs / peak, "phase_ms_per_step": ph,
                         "pressure_solve_frac": 3 * 12 * FT_B * N ** 3 / (ph["pressure_solve"] * 1e-3) / 1e9 / peak}
            del m
            torch.cuda.synchronize()
        return {"workload": "512^3 (P,P,P) Float64 WENO5 RK3 — the north_star target size", **res}

    def timed_model(make, ics, dt, warm, n, words, cells, label):
        """The same complete-path model in both arithmetic modes (exact: bit-identical to the reference's CPU arithmetic;
        contracted: FMA contraction + quotient-free WENO5 weights on the interior tiles)."""
        res = {"workload": label, "algorithmic_words_per_cell": words}
        for name, ex in (("exact", True), ("contracted", False)):
            m = make(ex)
            oc.set_model(m, **ics)
            oc.time_step(m, dt, n_steps=warm)
            ms = event_time(lambda: oc.time_step(m, dt, n_steps=n), torch.cuda.synchronize) / n
            gbs = words * FT_B * cells / (ms * 1e-3) / 1e9
            res[name] = {"ms_per_step": ms, "value": cells / (ms * 1e-3), "unit": UNIT, "steps": n, "whole_step_frac": gbs / peak,
                         "max_abs_divergence": m.max_abs_divergence()}
            del m
            torch.cuda.synchronize()
        return res

    def bounded_w(ics):
        ics["w"] = np.concatenate([ics["w"], ics["w"][:, :, :1]], axis=2)
        ics["w"][:, :, 0] = 0
        ics["w"][:, :, -1] = 0
        return ics

    def model_les():
        N = 256
        g = oc.RectilinearGrid(np.float64, size=(N, N, N), extent=(2 * np.pi,) * 3, topology=(oc.Periodic, oc.Periodic, oc.Bounded))
        make = lambda ex: oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), buoyancy=oc.SeawaterBuoyancy(), tracers=("T", "S"),   # noqa: E731
                                                 coriolis=oc.FPlane(1e-4), closure=oc.ScalarDiffusivity(1e-4, 1e-4), exact_arithmetic=ex)
        ics = bounded_w(smooth_ics(N, 1234))
        z = (np.arange(N) + 0.5) / N
        ics["T"] = 20 + 0.01 * z[None, None, :] + 1e-3 * ics["u"]
        ics["S"] = np.full((N, N, N), 35.0)
        # SURVEY.md §8d for F = 5 prognostic fields + pHY′: 124 words per cell and step
        return timed_model(make, ics, 0.2 * (2 * np.pi / N) / 2.0, 3, 5, 124, N ** 3,
                           "256^3 (P,P,B) Float64 WENO5 RK3, T + S, SeawaterBuoyancy, FPlane, ScalarDiffusivity")

    def model_c3():
        Nx, Nz = 512, 256
        g = oc.RectilinearGrid(np.float64, size=(Nx, Nx, Nz), x=(0, 2 * np.pi), y=(0, 2 * np.pi), z=tanh_faces(Nz),
                               topology=(oc.Periodic, oc.Periodic, oc.Bounded))
        make = lambda ex: oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), exact_arithmetic=ex)              # noqa: E731
        ics = bounded_w(smooth_ics(Nx, 1234, size=(Nx, Nx, Nz)))
        return timed_model(make, ics, 0.2 * (2 * np.pi / Nx) / 2.0, 2, 3, STEP_WORDS, Nx * Nx * Nz,
                           "configs[3]: 512x512x256 tanh-stretched Bounded z, WENO5 RK3, FourierTridiagonalPoissonSolver")

    guarded("step_512", step_512)
    guarded("solver_c2_f64", lambda: solver_target(oc, np.float64, (512, 512, 512), (0, 1)))
    guarded("solver_c2_f32", lambda: solver_target(oc, np.float32, (512, 512, 512), (0, 1)))
    guarded("solver_c3_f64", lambda: solver_target(oc, np.float64, (512, 512, 256), tanh_faces(256)))
    guarded("solver_c3_f32", lambda: solver_target(oc, np.float32, (512, 512, 256), tanh_faces(256)))
    guarded("model_c3_stretched", model_c3)
    guarded("model_les_256", model_les)
    return out


def slab_model(oc, D, world, local_rank, gsize, exact, extent_xy):
    arch = D.Distributed(D.Partition(y=world))
    return D.DistributedNonhydrostaticModel(arch, np.float64, size=gsize, x=(0, extent_xy[0]), y=(0, extent_xy[1]),
                                            z=(-2 * np.pi, 0), advection=oc.WENO(order=5), timestepper="RungeKutta3",
                                            device=local_rank, p2p=not os.environ.get("NHSTEP_NO_P2P"), exact_arithmetic=exact)


def multi_gpu_parity(oc, D, world, rank, local_rank, exact):
    """10 steps of a 64 x (32·N) x 32 problem on the N ranks AND on rank 0 alone (single-GPU model): the gathered u, v, w, p must
    agree to rel. L2 <= 1e-12 (the distributed FFT runs x, z, y instead of x, y, z, so agreement is to rounding) and the
    divergence must stay at rounding level.  The driver's test box has one GPU; this is where it sees the slab path checked."""
    import torch
    import torch.distributed as dist
    gsize = (64, 32 * world, 32)
    ics = smooth_ics(64, 4321, gsize)
    dt = 0.2 * (2 * np.pi / 64) / 2.0
    m = slab_model(oc, D, world, local_rank, gsize, exact, (2 * np.pi, 2 * np.pi * world / 2))
    m.set(**ics)
    m.time_step(dt, n_steps=10)
    got = {n: m.gather(n) for n in ("u", "v", "w", "pNHS")}
    div = m.max_abs_divergence()
    err = {}
    if rank == 0:
        g = oc.RectilinearGrid(np.float64, size=gsize, x=(0, 2 * np.pi), y=(0, 2 * np.pi * world / 2), z=(-2 * np.pi, 0),
                               topology=(oc.Periodic,) * 3, device=local_rank)
        s = oc.NonhydrostaticModel(g, advection=oc.WENO(order=5), timestepper="RungeKutta3", exact_arithmetic=exact)
        oc.set_model(s, **ics)
        oc.time_step(s, dt, n_steps=10)
        ref = {"u": s.velocities["u"].interior(), "v": s.velocities["v"].interior(), "w": s.velocities["w"].interior(),
               "pNHS": s.pressures.pNHS.interior()}
        for n in got:
            err[n] = float(np.linalg.norm((got[n] - ref[n]).ravel()) / np.linalg.norm(ref[n].ravel()))
        del s
    worst = torch.tensor([max(err.values()) if err else 0.0, div], device="cuda", dtype=torch.float64)
    dist.broadcast(worst, 0)
    del m
    torch.cuda.synchronize()
    ok = bool(worst[0].item() <= 1e-12 and worst[1].item() < 1e-10)
    return {"workload": f"{gsize[0]}x{gsize[1]}x{gsize[2]} (P,P,P) WENO5 RK3, 10 steps, {world} slabs vs one GPU", "vs_single_gpu": err,
            "max_div": div, "tolerance": 1e-12, "ok": ok}


def timed_slab_run(oc, D, world, rank, local_rank, gsize, extent_xy, exact, warmup, steps):
    import torch
    import torch.distributed as dist
    m = slab_model(oc, D, world, local_rank, gsize, exact, extent_xy)
    m.set_local(**smooth_ics(gsize[0], 1234, gsize, (m.j0, m.NyL)))
    N = gsize[2]
    dt = 0.2 * (2 * np.pi / N) / 2.0
    m.time_step(dt, n_steps=warmup)
    dist.barrier()
    ms = event_time(lambda: m.time_step(dt, n_steps=steps), torch.cuda.synchronize)
    t = torch.tensor([ms], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item()) / steps
    div = m.max_abs_divergence()
    del m
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    cells = gsize[0] * gsize[1] * gsize[2]
    peak, _ = peaks()
    gbs = STEP_WORDS * FT_B * cells / world / (ms * 1e-3) / 1e9
    return {"global_size": list(gsize), "ms_per_step": ms, "value": cells / (ms * 1e-3), "unit": UNIT, "steps": steps, "warmup": warmup,
            "whole_step_frac_per_gpu": gbs / peak, "max_abs_divergence": div}


def multi_gpu_targets(oc, D, world, rank, local_rank, exact):
    """512³ cells per GPU (the per-GPU size SURVEY.md §8e defines the ≥ 80 % weak-scaling target on; compare `value` / N with
    targets.step_512 of the N = 1 line) and BASELINE.json configs[4]: 1024³ global, strong scaling from 2 GPUs."""
    out = {}
    fx, fy = WORLD_FACTORS[world]

    def guarded(name, fn):
        try:
            out[name] = fn()
        except Exception as e:
            out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}

    # 512³ cells per GPU: the global grid grows in y, then x, then z, no direction beyond the 1024 points the transforms take
    dims, left = [512, 512, 512], world
    for e in (1, 0, 2):
        while left > 1 and dims[e] < 1024:
            dims[e] *= 2
            left //= 2
    guarded("c4_1024_cubed", lambda: {"scaling": "strong", **timed_slab_run(
        oc, D, world, rank, local_rank, (1024, 1024, 1024), (2 * np.pi, 2 * np.pi), exact, 2, 3)})
    if tuple(dims) == (1024, 1024, 1024):
        out["weak_512_per_gpu"] = {"scaling": "weak", "same_run_as": "c4_1024_cubed", **{k: v for k, v in out["c4_1024_cubed"].items() if k != "scaling"}}
    else:
        guarded("weak_512_per_gpu", lambda: {"scaling": "weak", **timed_slab_run(
            oc, D, world, rank, local_rank, tuple(dims), (2 * np.pi * dims[0] / 512, 2 * np.pi * dims[1] / 512), exact, 2, 5)})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-size", type=int, default=None, help="cells per side of the CPU sample (default: --size)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-targets", action="store_true", help="skip the `targets` / `parity` records (development)")
    # contracted: FMA contraction in the tendency sweep (what a GPU build of the reference computes; within the north_star
    # tolerance of the CPU reference); exact: bit-identical to the reference's CPU arithmetic.  Both are reported at N = 1.
    ap.add_argument("--arithmetic", choices=["contracted", "exact"], default="contracted")
    args = ap.parse_args()
    exact = args.arithmetic == "exact"
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import oceananigans_jl_b200 as oc
    if not torch.cuda.is_available() or oc.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        if world not in WORLD_FACTORS:
            raise SystemExit("--gpus must be 1, 2, 4, 8 or 16")

    N = args.size
    dt = 0.2 * (2 * np.pi / N) / 2.0
    parity = None
    if world == 1:
        gsize = (N, N, N)
        grid = oc.RectilinearGrid(np.float64, size=gsize, extent=(2 * np.pi,) * 3, topology=(oc.Periodic,) * 3, device=local_rank)
        model = oc.NonhydrostaticModel(grid, advection=oc.WENO(order=5), timestepper="RungeKutta3", exact_arithmetic=exact)
        oc.set_model(model, **smooth_ics(N, 1234))
        step = lambda n=1: oc.time_step(model, dt, n_steps=n)                              # noqa: E731
        profiled = lambda: oc.time_step_profiled(model, dt)                                # noqa: E731
    else:
        # weak scaling: N³ cells per GPU, the global grid grows in x and y at constant spacing; y-slabs over the ranks
        from oceananigans_jl_b200 import distributed as D
        if not args.no_targets:
            parity = multi_gpu_parity(oc, D, world, rank, local_rank, exact)
        fx, fy = WORLD_FACTORS[world]
        gsize = (N * fx, N * fy, N)
        model = slab_model(oc, D, world, local_rank, gsize, exact, (2 * np.pi * fx, 2 * np.pi * fy))
        model.set_local(**smooth_ics(N, 1234, gsize, (model.j0, model.NyL)))
        step = lambda n=1: model.time_step(dt, n_steps=n)                                  # noqa: E731
        profiled = lambda: model.time_step_profiled(dt)                                    # noqa: E731
    cells = gsize[0] * gsize[1] * gsize[2] // world                                        # per GPU

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = make_sampler(local_rank) if rank == 0 else None
    for _ in range(args.warmup):
        step()
    barrier()
    if sampler:
        sampler.wait_first_sample()
        sampler.mark()
    l0 = model.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    launches = model.launch_count() - l0
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = cells * world * args.steps / (ms * 1e-3)
    max_div = model.max_abs_divergence()
    assert max_div < 1e-9, f"the timed model lost incompressibility: max |div U| = {max_div}"

    # ---- phases: CUDA events on the model's stream between the phases of the very same time_step! (nh_time_step_profiled)
    phases = {"tendencies_substep": 0.0, "pressure_solve": 0.0, "pressure_correct": 0.0, "other": 0.0}
    nprof = max(1, min(args.steps, 5))
    for _ in range(nprof):
        for k, v in profiled().items():
            phases[k] = phases.get(k, 0.0) + v
    for k in phases:
        phases[k] /= nprof            # ms per step (3 substeps)
    # algorithmic bytes per step and phase (w = 8, F = 3): tendencies+substep A = 9w + 2·12w,
    # pressure solve B = 3·12w, pressure correction C = 3·7w; the dominant kernel is launched 3x per step
    alg = {"tendencies_substep": (9 + 2 * 12) * FT_B * cells, "pressure_solve": 3 * 12 * FT_B * cells,
           "pressure_correct": 3 * 7 * FT_B * cells}
    kernel_names = {"tendencies_substep": ("tendency_fast_kernel (tile sweep: fused WENO5 tendencies + RK3 substep)" if exact else
                                           "tendency_plane_kernel (TMA plane-marching sweep: fused WENO5 tendencies + RK3 substep)"),
                    "pressure_solve": "poisson (source term + transforms + spectral divide + real store)",
                    "pressure_correct": "pressure_correct_halo_kernel"}
    dom = max(alg, key=lambda k: phases[k])
    peak, peak_kind = peaks()
    achieved = alg[dom] / (phases[dom] * 1e-3) / 1e9
    step_bytes = STEP_WORDS * FT_B * cells
    roofline = {"bound": "hbm", "kernel": kernel_names[dom], "launches_per_step": 3,
                "algorithmic_bytes_per_launch": alg[dom] / 3, "ms_per_launch": phases[dom] / 3,
                "achieved": achieved, "peak": peak, "peak_source": peak_kind, "unit": "GB/s",
                "frac": achieved / peak, "traffic": measured_traffic(dom, N),
                "whole_step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms_per_step * 1e-3) / 1e9,
                               "frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak},
                "phase_ms_per_step": phases,
                "phase_frac": {k: alg[k] / (phases[k] * 1e-3) / 1e9 / peak for k in alg if phases[k] > 0}}
    if dom == "tendencies_substep":
        # the sweep is FP64-issue bound, not HBM bound: FP64 instructions per cell (ncu opcode mix,
        # profiles/) against 64 FP64 instr/clk/SM × 148 SMs at the sampled SM clock
        mhz = (clocks or {}).get("sm_mhz") or 1965.0
        fp64_peak = 64 * 148 * mhz * 1e6
        per_cell = FP64_INSTR_PER_CELL[args.arithmetic]
        fp64_rate = per_cell * cells / (phases[dom] / 3 * 1e-3)
        roofline["fp64_pipe"] = {"instr_per_cell": per_cell, "achieved_instr_per_s": fp64_rate, "peak_instr_per_s": fp64_peak,
                                 "frac": fp64_rate / fp64_peak}

    # ---- the other arithmetic mode, same workload (N = 1 only): a shorter timed run for the record ----
    other = None
    if world == 1:
        m2 = oc.NonhydrostaticModel(grid, advection=oc.WENO(order=5), timestepper="RungeKutta3", exact_arithmetic=not exact)
        oc.set_model(m2, **smooth_ics(N, 1234))
        n2 = max(1, min(args.steps, 20))
        oc.time_step(m2, dt, n_steps=max(3, args.warmup))
        torch.cuda.synchronize()
        e0.record()
        oc.time_step(m2, dt, n_steps=n2)
        e1.record()
        torch.cuda.synchronize()
        ms2 = e0.elapsed_time(e1) / n2
        other = {"arithmetic": "contracted" if exact else "exact", "ms_per_step": ms2, "value": cells / (ms2 * 1e-3),
                 "unit": UNIT, "steps": n2}
        del m2

    # ---- e2e: HOST buffers through the C ABI; H2D of u,v,w and D2H of u,v,w,p inside the timed region ----
    from oceananigans_jl_b200.model import pinned_empty
    shape = model.velocities["u"].size
    host_np = {n: pinned_empty(shape) for n in "uvw"}            # the step's inputs, page-locked
    host_out = {n: pinned_empty(shape) for n in ("u", "v", "w", "p")}
    for n in "uvw":
        model.velocities[n].interior(out=host_np[n])
    e2e_steps = max(1, min(args.steps, 10))
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        if world == 1:
            for n in "uvw":
                model.velocities[n].set(host_np[n])
            oc.fill_halo_regions(model.prognostic_fields())
        else:
            model.set_local(enforce_incompressibility=False, **host_np)
        step()
        for n in "uvw":
            model.velocities[n].interior(out=host_out[n])
        model.pressures.pNHS.interior(out=host_out["p"])
        for n in "uvw":                                          # next step's inputs = this step's result (no copy)
            host_np[n], host_out[n] = host_out[n], host_np[n]
    barrier()
    e2e_t = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_t], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_t = float(t.item())
    e2e = {"value": cells * world * e2e_steps / e2e_t, "unit": UNIT, "h2d_bytes_per_step": 3 * cells * FT_B * world,
           "d2h_bytes_per_step": 4 * cells * FT_B * world, "steps": e2e_steps}

    # ---- the rest of the contract's configurations, in the same process ----
    targets = None
    if not args.no_targets:
        del model, host_np, host_out
        torch.cuda.synchronize()
        torch.cuda.empty_cache()
        if world == 1:
            targets = single_gpu_targets(oc, args, exact)
        else:
            targets = multi_gpu_targets(oc, D, world, rank, local_rank, exact)

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            rs = args.ref_size or args.size
            rate, sps, cores, done = cpu_oracle_rate(rs, 10, 1, budget_s=25.0)
            cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": CPU_SAMPLE.format(n=rs, k=done, c=cores, s=sps)}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(args, world), "other_arithmetic": other,
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
                "max_abs_divergence": max_div}
        if parity is not None:
            line["parity"] = parity
        if targets is not None:
            line["targets"] = targets
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

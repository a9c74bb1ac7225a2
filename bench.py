#!/usr/bin/env python
"""bench.py — cell-updates/s of the NonhydrostaticModel WENO5 RK3 time step (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--size 256] [--impl reference]

A "step" is one time_step!(model, Δt) (3 RK3 substeps incl. 3 pressure solves) on a triply periodic Float64 grid;
one cell-update = one grid cell advanced by one step.  Prints ONE JSON line (see the task contract): `value` is the
device-resident throughput (CUDA events), `e2e` the same step driven through the C ABI with HOST buffers (u, v, w copied
host→device before and u, v, w, p device→host after every step), `roofline` the dominant phase against measured HBM
bandwidth, `cpu_baseline` the compiled C / OpenMP restatement of the reference's CPU path (oracle/cpu_step.c) on a bounded sample,
all host threads.
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
# FP64-pipe instructions per cell of tendency_fast_kernel (ncu opcode mix, profiles/r1_tendency_sass_mix.md)
FP64_INSTR_PER_CELL = {"exact": 1290.0, "contracted": 750.0}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def measured_traffic(phase, N):
    """DRAM bytes per launch of the phase's kernel(s) from the committed ncu capture, if it was taken at this size."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
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


def cpu_oracle_rate(N, steps, warmup=1):
    """The compiled C / OpenMP restatement of the reference's CPU() path (oracle/cpu_step.c: one pass per reference
    kernel, every face flux evaluated twice, full complex FFT — pinned against the NumPy oracle in tests/test_cpu_step.py)
    on an N³ sample of the same workload, on all host threads OpenMP gives it.  Returns (rate, s/step, threads)."""
    from oracle.cpu_step import CpuStepModel
    CpuStepModel.use_all_cores()
    m = CpuStepModel((N, N, N), (2 * np.pi,) * 3)
    ics = smooth_ics(N, 1234)
    m.set(ics["u"], ics["v"], ics["w"])
    dt = 0.2 * (2 * np.pi / N) / 2.0
    m.time_step(dt, max(1, warmup))
    t0 = time.perf_counter()
    m.time_step(dt, steps)
    t = time.perf_counter() - t0
    return N ** 3 * steps / t, t / steps, m.threads()


CPU_SAMPLE = ("oracle/cpu_step.c (C / OpenMP restatement of the reference CPU() path; Julia is absent on this box), "
              "{n}^3 cells x {k} steps on {c} threads, {s:.3f} s/step")


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


def run_reference(args, rank):
    if rank != 0:
        return
    N = args.ref_size
    rate, sps, cores = cpu_oracle_rate(N, max(1, args.steps), max(1, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"3D triply-periodic NonhydrostaticModel {args.size}^3 Float64 WENO5 RK3 FFTBasedPoissonSolver",
                       "sample": f"{N}^3 cells per step"},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": CPU_SAMPLE.format(n=N, k=args.steps, c=cores, s=sps)},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--size", type=int, default=256)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-size", type=int, default=128)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    # contracted: FMA contraction in the tendency sweep (what a GPU build of the reference computes; within the north_star
    # tolerance of the CPU reference); exact: bit-identical to the reference's CPU arithmetic.  Both are reported at N = 1.
    ap.add_argument("--arithmetic", choices=["contracted", "exact"], default="contracted")
    args = ap.parse_args()
    exact = args.arithmetic == "exact"
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    import oceananigans_jl_b200 as oc
    if not torch.cuda.is_available() or oc.device_count() < 1:
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    N = args.size
    dt = 0.2 * (2 * np.pi / N) / 2.0
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
        factors = {2: (1, 2), 4: (2, 2), 8: (2, 4), 16: (4, 4)}
        if world not in factors:
            raise SystemExit("--gpus must be 1, 2, 4, 8 or 16")
        fx, fy = factors[world]
        gsize = (N * fx, N * fy, N)
        arch = D.Distributed(D.Partition(y=world))
        model = D.DistributedNonhydrostaticModel(arch, np.float64, size=gsize, x=(0, 2 * np.pi * fx), y=(0, 2 * np.pi * fy),
                                                 z=(-2 * np.pi, 0), advection=oc.WENO(order=5), timestepper="RungeKutta3",
                                                 device=local_rank, p2p=not os.environ.get("NHSTEP_NO_P2P"),
                                                 exact_arithmetic=exact)
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
    assert np.isfinite(model.max_abs_divergence()), "model blew up"

    # ---- phases (per-kernel-group CUDA-event times; same step composed from the individual ABI calls) ----
    FT_B = 8
    # CUDA events on the model's stream between the phases of the very same time_step! (nh_time_step_profiled)
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
    kernel_names = {"tendencies_substep": "tendency_fast_kernel (fused WENO5 tendencies + RK3 substep)",
                    "pressure_solve": "poisson (source term + transforms + spectral divide + real store)",
                    "pressure_correct": "pressure_correct_halo_kernel"}
    dom = max(alg, key=lambda k: phases[k])
    peak, peak_kind = peaks()
    achieved = alg[dom] / (phases[dom] * 1e-3) / 1e9
    step_bytes = 90 * FT_B * cells
    roofline = {"bound": "hbm", "kernel": kernel_names[dom], "launches_per_step": 3,
                "algorithmic_bytes_per_launch": alg[dom] / 3, "ms_per_launch": phases[dom] / 3,
                "achieved": achieved, "peak": peak, "peak_source": peak_kind, "unit": "GB/s",
                "frac": achieved / peak, "traffic": measured_traffic(dom, N),
                "whole_step": {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms_per_step * 1e-3) / 1e9,
                               "frac": step_bytes / (ms_per_step * 1e-3) / 1e9 / peak},
                "phase_ms_per_step": phases}
    if dom == "tendencies_substep":
        # the sweep is FP64-issue bound, not HBM bound: FP64 instructions per cell (ncu opcode mix,
        # profiles/r1_tendency_sass_mix.md) against 64 FP64 instr/clk/SM × 148 SMs at the sampled SM clock
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

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            rate, sps, cores = cpu_oracle_rate(args.ref_size, 10)
            cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": CPU_SAMPLE.format(n=args.ref_size, k=10, c=cores, s=sps)}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": f"3D triply-periodic NonhydrostaticModel {N}^3 Float64 WENO5 RK3 FFTBasedPoissonSolver"
                           + (f" per GPU (global {gsize[0]}x{gsize[1]}x{gsize[2]})" if world > 1 else ""),
                           "arithmetic": ("exact: bit-identical to the reference's CPU arithmetic (no FMA contraction)" if exact else
                                          "contracted: FMA contraction + regrouped WENO5 weights in the tendency sweep, "
                                          "rel. L2 <= 1e-12 vs the oracle after 100 steps (tests/test_gpu_parity.py)"),
                           "other_arithmetic": other,
                           "l2_policy": "working set (>= 10 fields x %.0f MB) exceeds the 126 MB L2" % (cells * FT_B / 1e6),
                           "parallelism": "single GPU" if world == 1 else
                           (f"y-slabs over {world} GPUs: kernels store halo rows and FFT transpose blocks into the peers' "
                            "buffers over NVLink (CUDA IPC), 5 one-element NCCL all-reduces per stage as barriers"
                            if getattr(model, "p2p", False) else
                            f"y-slabs over {world} GPUs: NCCL halo exchange + 2 all-to-all transposes per pressure solve")},
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

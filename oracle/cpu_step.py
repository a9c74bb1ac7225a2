"""ctypes binding of oracle/cpu_step.c — the compiled C / OpenMP restatement of the reference's CPU() path for the
benchmark workload (regular triply-periodic grid, WENO5, RK3, FFT pressure solve, Float64).  Test infrastructure and
timed CPU baseline only: imported by tests/, bench.py's cpu_baseline / --impl reference legs and nothing else.
Coefficients, spacings and eigenvalues are taken from the NumPy oracle's objects so both share the same bits."""
import ctypes
import os
import subprocess

import numpy as np

from .grid import RectilinearGrid, Periodic
from .advection import WENO
from .solvers import poisson_eigenvalues
from .grid import round_fraction
from fractions import Fraction

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libcpustep.so")
        src = os.path.join(_HERE, "cpu_step.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.run(["make", "-C", _HERE, "libcpustep.so"], check=True, capture_output=True)
        L = ctypes.CDLL(so)
        dp, ip = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
        L.cpu_model_create.restype = ctypes.c_void_p
        L.cpu_model_create.argtypes = [ip, dp, dp, dp, dp, dp, dp, dp, dp, ctypes.c_double, dp]
        L.cpu_model_destroy.argtypes = [ctypes.c_void_p]
        L.cpu_model_get.argtypes = [ctypes.c_void_p, ctypes.c_int, dp]
        L.cpu_model_set.argtypes = [ctypes.c_void_p, dp, dp, dp, ctypes.c_int]
        L.cpu_model_compute_tendencies.argtypes = [ctypes.c_void_p]
        L.cpu_model_time_step.argtypes = [ctypes.c_void_p, ctypes.c_double]
        L.cpu_threads.restype = ctypes.c_int
        L.cpu_set_threads.argtypes = [ctypes.c_int]
        _LIB = L
    return _LIB


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


FIELDS = {"u": 0, "v": 1, "w": 2, "p": 3, "Gu": 4, "Gv": 5, "Gw": 6, "Gmu": 7, "Gmv": 8, "Gmw": 9}


class CpuStepModel:
    """NonhydrostaticModel(grid = RectilinearGrid(size, extent, topology = (Periodic, Periodic, Periodic)),
    advection = WENO(order = 5), timestepper = :RungeKutta3) on the host cores."""

    def __init__(self, size, extent):
        FT = np.float64
        g = RectilinearGrid(FT, size=tuple(size), extent=tuple(extent), topology=(Periodic,) * 3)
        s = WENO(FT, 5)
        self.N = tuple(int(n) for n in size)
        N = np.array(self.N, dtype=np.int32)
        d = np.array([float(g.coords[a].dc) for a in range(3)])
        L = (g.Lx, g.Ly, g.Lz)
        lam = [np.ascontiguousarray(poisson_eigenvalues(self.N[a], L[a], Periodic), dtype=np.float64) for a in range(3)]
        csym = np.array(s.advecting_velocity_scheme.sym_coeff, dtype=np.float64)
        cw = np.array(s.coeff, dtype=np.float64).reshape(9)
        cstar = np.array(s.cstar, dtype=np.float64)
        sm = np.array(s.smooth, dtype=np.float64).reshape(18)
        rk3 = np.array([round_fraction(Fraction(*f), FT) for f in ((8, 15), (5, 12), (3, 4), (-17, 60), (-5, 12))])
        self._h = lib().cpu_model_create(N.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), _dp(d), _dp(lam[0]), _dp(lam[1]),
                                         _dp(lam[2]), _dp(csym), _dp(cw), _dp(cstar), _dp(sm), float(s.eps), _dp(rk3))
        if not self._h:
            raise ValueError("cpu_step: sizes must be powers of two")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().cpu_model_destroy(self._h)
            self._h = None

    def set(self, u, v, w, enforce_incompressibility=True):
        arrs = [np.ascontiguousarray(np.asarray(a, dtype=np.float64).transpose(2, 1, 0)) for a in (u, v, w)]
        lib().cpu_model_set(self._h, _dp(arrs[0]), _dp(arrs[1]), _dp(arrs[2]), int(enforce_incompressibility))

    def compute_tendencies(self):
        lib().cpu_model_compute_tendencies(self._h)

    def time_step(self, dt, n_steps=1):
        for _ in range(n_steps):
            lib().cpu_model_time_step(self._h, float(dt))

    def get(self, name):
        out = np.empty(self.N[::-1], dtype=np.float64)
        lib().cpu_model_get(self._h, FIELDS[name], _dp(out))
        return out.transpose(2, 1, 0)

    @staticmethod
    def threads():
        return int(lib().cpu_threads())

    @staticmethod
    def use_all_cores():
        """All cores this process may run on, whatever OMP_NUM_THREADS says (torchrun sets it to 1 for its ranks)."""
        n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        lib().cpu_set_threads(n)
        return n

"""ORACLE (test infrastructure): loader of the C/OpenMP port `oracle/mg_oracle_c.c`.

The port exports a subset of the C ABI of include/mgengine.h (scalar real fields, Dirichlet
boundaries, CG coarse solver, the FAS stages), so the product's ctypes `Workspace` class can drive
it when it is handed this library.  Only tests/ and bench.py's CPU legs import this module; nothing
under evostencils_b200/ or scripts/ does."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from evostencils_b200 import engine, fas

ORACLE_DIR = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(ORACLE_DIR, "_build", "libmgoracle.so")
_lib = None


def build(march=None, force=False):
    cmd = ["make", "-s", "-C", ORACLE_DIR]
    if force:
        cmd.append("-B")
    if march:
        cmd.append(f"MARCH={march}")
    try:
        subprocess.run(cmd, check=True, capture_output=True, timeout=180)
    except Exception:
        if not os.path.exists(LIB):
            raise


def load(march=None, force=False):
    global _lib
    if _lib is None or force:
        build(march, force)
        lib = engine.bind(C.CDLL(LIB), strict=False)
        lib.mg_oracle_set_threads.argtypes = [C.c_int]
        lib.mg_oracle_threads.restype = C.c_int
        _lib = lib
    return _lib


def set_threads(n: int) -> int:
    """Force the OpenMP team size (whatever OMP_NUM_THREADS says) -> threads a parallel region gets."""
    lib = load()
    lib.mg_oracle_set_threads(int(n))
    return int(lib.mg_oracle_threads())


def workspace(dimension, min_level, max_level, n_fields=1, dtype="f64"):
    return engine.Workspace(dimension, min_level, max_level, n_fields, dtype, 0, lib=load())


class FasSolver(fas.FasSolver):
    """The host-side FAS driver of the product on the C port's stages (CPU figure / checker)."""

    def _make_workspace(self, device):
        self.use_graph = False                      # no stream, no graph: the port runs every statement at once
        return workspace(2, self.min_level, self.max_level)

    def _prolong_add(self, op):
        from evostencils_b200 import lowering
        self.ws.exec_ops([lowering.Op("prolong_add", op.level, omega=op.omega, src="sol", dst="sol")])

    def _timer(self, stop):
        import time
        now = time.perf_counter()
        if not stop:
            self._t0 = now
            return None
        return (now - self._t0) * 1e3

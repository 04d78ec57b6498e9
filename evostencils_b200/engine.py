"""ctypes binding of the C-ABI engine library (include/mgengine.h, csrc/mg_engine.cu).

This is the host side of the drop-in boundary: it turns a lowered `Program`
(lowering.py) into `mg_op` records plus smoother tables and drives one workspace.
There is no CPU fallback: if the shared library is missing or no CUDA device is
present, construction raises.
"""
from __future__ import annotations

import ctypes as C
import itertools
import os
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import lowering

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmgengine.so")

MG_F64, MG_C128 = 0, 1
MG_SOL, MG_RHS, MG_RES, MG_COR = 0, 1, 2, 3
MG_BC_DIRICHLET, MG_BC_HELMHOLTZ_ROBIN = 0, 1
MG_CGS_CG, MG_CGS_BICGSTAB = 0, 1
MG_VC_NONE, MG_VC_EXP_KAPPA = 0, 1
(MG_OP_SMOOTH, MG_OP_RESIDUAL, MG_OP_RESTRICT, MG_OP_RESIDUAL_RESTRICT, MG_OP_ZERO, MG_OP_PROLONG_ADD,
 MG_OP_PROLONG, MG_OP_CGS, MG_OP_KRYLOV) = range(9)
MG_KRYLOV_METHOD = {"ConjugateGradient": 0, "BiCGStab": 1}

_FIELD = {"sol": MG_SOL, "rhs": MG_RHS, "res": MG_RES, "cor": MG_COR, "": MG_SOL}

EXPORTED_SYMBOLS = [
    "mg_create", "mg_create_slab", "mg_field_pointer", "mg_destroy", "mg_last_error", "mg_launch_count", "mg_set_operator", "mg_set_boundary",
    "mg_set_variable_coefficient", "mg_set_coarse_solver", "mg_add_smoother", "mg_clear_smoothers", "mg_upload", "mg_download",
    "mg_upload_device", "mg_save_initial_guess", "mg_restore_initial_guess", "mg_exec_ops", "mg_enqueue_ops", "mg_set_stream", "mg_build_cycle",
    "mg_free_cycles", "mg_capture_begin", "mg_capture_end", "mg_graph_launch", "mg_run_cycles", "mg_residual_norm", "mg_solve", "mg_set_time_limit", "mg_set_outer_operator",
    "mg_solve_bicgstab", "mg_synchronize",
    "mg_time_ops", "mg_profile_ops", "mg_mark", "mg_elapsed", "mg_flush_l2", "mg_set_fast_kernels",
    "mg_fas_smooth", "mg_fas_residual", "mg_fas_restrict", "mg_fas_subtract_approximation", "mg_fas_residual_norm",
    "mg_residual_norm_noalloc",
    "mg_dd_init", "mg_dd_describe", "mg_dd_attach", "mg_dd_exchange", "mg_dd_build_cycle", "mg_dd_prepare", "mg_dd_launch",
    "mg_dd_norm", "mg_dd_read_norm", "mg_dd_error",
]


class MgOp(C.Structure):
    _fields_ = [("kind", C.c_int32), ("level", C.c_int32), ("omega", C.c_double), ("src", C.c_int32),
                ("smoother", C.c_int32)]


MG_ERR_INVALID, MG_ERR_CUDA, MG_ERR_STATE, MG_ERR_NOMEM, MG_ERR_TIMEOUT = -1, -2, -3, -4, -5


class EngineError(RuntimeError):
    """`code` = the mg_status of the failing call (0: raised by the host side).  `fatal`: the engine
    itself is broken (CUDA failure, out of memory, library missing) -- everything else is a property
    of the individual being evaluated and maps to the reference's `(infinity,)*3` sentinel
    (code_generation/exastencils.py:424-431)."""

    def __init__(self, message, code=0, fatal=False):
        super().__init__(message)
        self.code = code
        self.fatal = fatal or code in (MG_ERR_CUDA, MG_ERR_NOMEM)


_lib = None


def _signatures():
    vp, i32, dbl = C.c_void_p, C.c_int, C.c_double
    pi32, pdbl, pint, pflt = C.POINTER(C.c_int32), C.POINTER(dbl), C.POINTER(C.c_int), C.POINTER(C.c_float)
    ops = C.POINTER(MgOp)
    return {
        "mg_create": [C.POINTER(vp), i32, i32, i32, i32, i32, i32],
        "mg_create_slab": [C.POINTER(vp), i32, i32, i32, i32, i32, pi32, pi32],
        "mg_field_pointer": [vp, i32, i32, i32, C.POINTER(vp), pi32, pi32, pi32, pi32],
        "mg_destroy": [vp],
        "mg_last_error": [],
        "mg_launch_count": [vp, C.POINTER(C.c_int64)],
        "mg_set_operator": [vp, i32, i32, i32, i32, pi32, pdbl],
        "mg_set_boundary": [vp, i32, dbl],
        "mg_set_variable_coefficient": [vp, i32, dbl],
        "mg_set_coarse_solver": [vp, i32, i32, dbl],
        "mg_add_smoother": [vp, i32, i32, i32, pi32, i32, pi32, pi32, i32, C.POINTER(C.c_uint32), pdbl, pint],
        "mg_clear_smoothers": [vp],
        "mg_capture_begin": [vp],
        "mg_capture_end": [vp, pint],
        "mg_graph_launch": [vp, i32, i32],
        "mg_upload": [vp, i32, i32, i32, vp],
        "mg_download": [vp, i32, i32, i32, vp],
        "mg_upload_device": [vp, i32, i32, i32, vp],
        "mg_save_initial_guess": [vp],
        "mg_restore_initial_guess": [vp],
        "mg_exec_ops": [vp, ops, i32],
        "mg_enqueue_ops": [vp, ops, i32],
        "mg_set_stream": [vp, vp],
        "mg_build_cycle": [vp, ops, i32, pint],
        "mg_free_cycles": [vp],
        "mg_run_cycles": [vp, i32, i32, pflt],
        "mg_residual_norm": [vp, i32, pdbl],
        "mg_solve": [vp, i32, dbl, i32, pdbl, pint, pint, pflt],
        "mg_set_time_limit": [vp, dbl],
        "mg_set_outer_operator": [vp, i32, pi32, pdbl],
        "mg_solve_bicgstab": [vp, i32, dbl, i32, pdbl, pint, pint, pflt],
        "mg_synchronize": [vp],
        "mg_time_ops": [vp, ops, i32, i32, pflt],
        "mg_profile_ops": [vp, ops, i32, pflt],
        "mg_mark": [vp, i32],
        "mg_elapsed": [vp, pflt],
        "mg_flush_l2": [vp, C.c_size_t],
        "mg_set_fast_kernels": [vp, i32],
        "mg_fas_smooth": [vp, i32, dbl, dbl, i32, i32, i32],
        "mg_fas_residual": [vp, i32, dbl],
        "mg_fas_restrict": [vp, i32, dbl],
        "mg_fas_subtract_approximation": [vp, i32],
        "mg_fas_residual_norm": [vp, i32, dbl, pdbl],
        "mg_residual_norm_noalloc": [i32, i32, i32, i32, pi32, pdbl, vp, vp, vp, vp],
        "mg_dd_init": [vp, i32, i32, vp],
        "mg_dd_describe": [vp, i32, C.POINTER(C.c_uint64), vp],
        "mg_dd_attach": [vp, i32, i32, C.POINTER(C.c_uint64), vp],
        "mg_dd_exchange": [vp, i32, i32, i32],
        "mg_dd_build_cycle": [vp, ops, i32, pint],
        "mg_dd_prepare": [vp, i32],
        "mg_dd_launch": [vp, i32, i32],
        "mg_dd_norm": [vp],
        "mg_dd_read_norm": [vp, pdbl],
        "mg_dd_error": [vp, pint],
    }


def bind(lib, strict=True):
    """Attach the argument types of include/mgengine.h to a loaded library.  `strict`: every
    declared entry point must be exported (the engine); otherwise absent ones are skipped (a library
    that implements a subset of the ABI)."""
    for name, argtypes in _signatures().items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if strict:
                raise EngineError(f"the engine library does not export {name}", fatal=True) from None
            continue
        fn.argtypes = argtypes
    lib.mg_last_error.restype = C.c_char_p
    return lib


def load_library():
    """dlopen the engine; raises (loudly) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("MG_ENGINE_LIB", LIB_PATH)      # another build of the same sources (e.g. -DMG_TAIL_TIMING)
    if not os.path.exists(path):
        raise EngineError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the engine has no CPU fallback)", fatal=True)
    _lib = bind(C.CDLL(path))
    return _lib


def residual_norm_noalloc(dimension, n, pitch, entries, u_device, rhs_device, out_device, cuda_stream=0):
    """mg_residual_norm_noalloc (include/mgengine.h): the residual norm of one scalar fp64 field, enqueued
    without any allocation; `entries` = [(offset, value), ...], the pointers are device addresses."""
    lib = load_library()
    k = len(entries)
    offs = (C.c_int32 * max(1, k * dimension))(*[int(c) for off, _ in entries for c in off])
    vals = (C.c_double * max(1, k))(*[float(v) for _, v in entries])
    rc = lib.mg_residual_norm_noalloc(dimension, n, pitch, k, offs, vals, C.c_void_p(u_device), C.c_void_p(rhs_device),
                                      C.c_void_p(out_device), C.c_void_p(cuda_stream or None))
    if rc != 0:
        raise EngineError(f"mgengine error {rc}: {lib.mg_last_error().decode()}", code=rc)


# ----------------------------------------------------------------------------------------
# smoother tables: W = omega * inv(L) for every boundary class of a tile
# ----------------------------------------------------------------------------------------

def _omega_inverse(matrix: np.ndarray, omega: float) -> np.ndarray:
    m = matrix.shape[0]
    if m == 1:
        return np.array([[omega / matrix[0, 0]]], dtype=matrix.dtype)
    if m == 2:
        a, b, c, d = matrix[0, 0], matrix[0, 1], matrix[1, 0], matrix[1, 1]
        det = a * d - b * c
        return np.array([[omega * (d / det), omega * (-b / det)],
                         [omega * (-c / det), omega * (a / det)]], dtype=matrix.dtype)
    return omega * np.linalg.inv(matrix)


def smoother_tables(stmt: lowering.LocalSolve, dimension: int, dtype):
    """-> (masks uint32[k], W[k][m][m]) for the validity masks that occur on this level."""
    n = 2 ** stmt.level + 1
    m = len(stmt.slots)
    matrix = np.array(stmt.matrix, dtype=dtype)
    per_dim = []
    for ax in range(dimension):
        b = stmt.period[ax]
        classes = set()
        for t in range((n - 1) // b + 1):
            classes.add(tuple(1 <= t * b + p <= n - 2 for p in range(b)))
        per_dim.append(sorted(classes))
    tables: Dict[int, np.ndarray] = {}
    for combo in itertools.product(*per_dim):
        valid = [all(combo[ax][pos[ax]] for ax in range(dimension)) for _, pos in stmt.slots]
        mask = sum(1 << s for s in range(m) if valid[s])
        if mask == 0 or mask in tables:
            continue
        idx = [s for s in range(m) if valid[s]]
        w = np.zeros((m, m), dtype=dtype)
        w[np.ix_(idx, idx)] = _omega_inverse(matrix[np.ix_(idx, idx)], stmt.omega)
        tables[mask] = w
    full = (1 << m) - 1
    if full not in tables:      # grid smaller than a tile: the kernel still needs the slot
        tables[full] = _omega_inverse(matrix, stmt.omega)
    masks = np.array(sorted(tables), dtype=np.uint32)
    w_all = np.ascontiguousarray(np.stack([tables[int(k)] for k in masks]))
    return masks, w_all


class Workspace:
    """One level hierarchy on one GPU.  Reused across individuals of the same problem."""

    def __init__(self, dimension, min_level, max_level, n_fields, dtype="f64", device=0, lib=None, slab=None):
        """`slab` = {level: (first owned plane, last owned plane)} makes a z-slab of a domain-decomposed
        3D grid (mg_create_slab); host arrays are then [nz][n][n] with one ghost plane on either side."""
        self.lib = lib or load_library()
        self.dimension, self.min_level, self.max_level = dimension, min_level, max_level
        self.n_fields = n_fields
        self.dtype = dtype
        self.np_dtype = np.complex128 if dtype == "c128" else np.float64
        self.device = device
        self.slab = dict(slab) if slab else None
        self._h = C.c_void_p()
        code = MG_C128 if dtype == "c128" else MG_F64
        if self.slab:
            if dimension != 3:
                raise EngineError("z-slabs exist for 3D grids only")
            count = max_level - min_level + 1
            zb = (C.c_int32 * count)(*[self.slab[level][0] for level in range(min_level, max_level + 1)])
            ze = (C.c_int32 * count)(*[self.slab[level][1] for level in range(min_level, max_level + 1)])
            self._check(self.lib.mg_create_slab(C.byref(self._h), min_level, max_level, n_fields, code, device, zb, ze))
        else:
            self._check(self.lib.mg_create(C.byref(self._h), dimension, min_level, max_level, n_fields, code, device))
        self._smoother_cache: Dict[tuple, int] = {}
        self._operators: Dict[int, list] = {}
        self._variable = None

    # -- plumbing ------------------------------------------------------------------------
    def _check(self, rc):
        if rc != 0:
            raise EngineError(f"mgengine error {rc}: {self.lib.mg_last_error().decode()}", code=rc)

    def close(self):
        if self._h:
            self.lib.mg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- problem definition ----------------------------------------------------------------
    def set_operator(self, level, table):
        """`table` = F x F canonical stencils [(offset, value), ...] (lowering.operator_table)."""
        if self._operators.get(level) == table:
            return
        for i, row in enumerate(table):
            for j, entries in enumerate(row):
                n = len(entries)
                offs = (C.c_int32 * max(1, n * self.dimension))()
                vals = (C.c_double * max(1, n * (2 if self.dtype == "c128" else 1)))()
                for k, (off, val) in enumerate(entries):
                    for c in range(self.dimension):
                        offs[k * self.dimension + c] = off[c]
                    if self.dtype == "c128":
                        v = complex(val)
                        vals[2 * k], vals[2 * k + 1] = v.real, v.imag
                    else:
                        vals[k] = float(val)
                self._check(self.lib.mg_set_operator(self._h, level, i, j, n, offs, vals))
        self._operators[level] = table

    def set_outer_operator(self, entries):
        """Unshifted operator A of the finest level for the outer BiCGStab: [(offset, value), ...]."""
        n = len(entries)
        offs = (C.c_int32 * (n * self.dimension))()
        vals = (C.c_double * (n * (2 if self.dtype == "c128" else 1)))()
        for k, (off, val) in enumerate(entries):
            for c in range(self.dimension):
                offs[k * self.dimension + c] = off[c]
            if self.dtype == "c128":
                v = complex(val)
                vals[2 * k], vals[2 * k + 1] = v.real, v.imag
            else:
                vals[k] = float(val)
        self._check(self.lib.mg_set_outer_operator(self._h, n, offs, vals))

    def set_boundary(self, kind: str, k: float = 0.0):
        self._check(self.lib.mg_set_boundary(
            self._h, MG_BC_HELMHOLTZ_ROBIN if kind == "helmholtz_robin" else MG_BC_DIRICHLET, float(k)))

    def set_variable_coefficient(self, spec):
        """`spec` = {"kind": "exp_kappa", "kappa": 10.0} (problems.ProblemSpec.variable_coefficient) or None:
        the operator of every level becomes -div(c grad u) with c evaluated per node on the device."""
        spec = dict(spec) if spec else None
        if spec == self._variable:
            return
        if spec is None:
            self._check(self.lib.mg_set_variable_coefficient(self._h, MG_VC_NONE, 0.0))
        else:
            if spec.get("kind") != "exp_kappa":
                raise EngineError(f"unknown variable coefficient {spec!r}")
            self.clear_programs(smoothers=True)
            self._check(self.lib.mg_set_variable_coefficient(self._h, MG_VC_EXP_KAPPA, float(spec["kappa"])))
        self._variable = spec

    def set_coarse_solver(self, kind: str, max_iterations: int, reduction: float):
        self._check(self.lib.mg_set_coarse_solver(
            self._h, MG_CGS_CG if kind == "CG" else MG_CGS_BICGSTAB, int(max_iterations), float(reduction)))

    def smoother_handle(self, stmt: lowering.LocalSolve) -> int:
        key = (stmt.key(), stmt.matrix)
        if key in self._smoother_cache:
            return self._smoother_cache[key]
        m = len(stmt.slots)
        if self._variable:
            # the centre value depends on the node: the table holds omega alone and the kernels divide by
            # the node's a00 (mg_set_variable_coefficient)
            if m != 1 or any(p != 1 for p in stmt.period):
                raise EngineError("block smoothers are not available for variable-coefficient operators")
            masks, w = np.array([1], dtype=np.uint32), np.array([[[stmt.omega]]], dtype=self.np_dtype)
        else:
            masks, w = smoother_tables(stmt, self.dimension, self.np_dtype)
        period = (C.c_int32 * 3)(*(list(stmt.period) + [1] * (3 - len(stmt.period))))
        slot_field = (C.c_int32 * m)(*[f for f, _ in stmt.slots])
        slot_pos = (C.c_int32 * (m * self.dimension))(*[c for _, pos in stmt.slots for c in pos])
        handle = C.c_int()
        wflat = np.ascontiguousarray(w).view(np.float64).reshape(-1)
        self._check(self.lib.mg_add_smoother(
            self._h, stmt.level, int(stmt.jacobi), int(stmt.colored), period, m, slot_field, slot_pos,
            len(masks), masks.ctypes.data_as(C.POINTER(C.c_uint32)), wflat.ctypes.data_as(C.POINTER(C.c_double)),
            C.byref(handle)))
        self._smoother_cache[key] = handle.value
        return handle.value

    MAX_CACHED_SMOOTHERS = 2048

    def clear_programs(self, smoothers=False):
        """Drop the captured cycles (between individuals).  Smoother tables are keyed by their
        content (level, weight, partitioning, tile shape, local matrix), so they stay valid for the
        next individual and are kept unless `smoothers` is set, the operators changed
        (set_operator) or the cache has grown beyond MAX_CACHED_SMOOTHERS tables."""
        if smoothers or len(self._smoother_cache) > self.MAX_CACHED_SMOOTHERS:
            self._check(self.lib.mg_clear_smoothers(self._h))
            self._smoother_cache.clear()
        else:
            self._check(self.lib.mg_free_cycles(self._h))

    # -- op encoding -------------------------------------------------------------------------
    def encode(self, ops: Sequence[object], default_cycle=None) -> "C.Array":
        flat: List[MgOp] = []

        def add(kind, level, omega=1.0, src=MG_SOL, smoother=-1):
            flat.append(MgOp(kind, level, float(omega), src, smoother))

        def walk(seq):
            for op in seq:
                k = op.kind
                if k == "smooth":
                    add(MG_OP_SMOOTH, op.level, op.omega, int(getattr(op, "colour_pass", 0)), self.smoother_handle(op))
                elif k == "residual":
                    add(MG_OP_RESIDUAL, op.level)
                elif k == "restrict":
                    add(MG_OP_RESTRICT, op.level, 1.0, _FIELD[op.src or "res"])
                elif k == "residual_restrict":
                    add(MG_OP_RESIDUAL_RESTRICT, op.level)
                elif k == "zero":
                    add(MG_OP_ZERO, op.level)
                elif k == "prolong_add":
                    add(MG_OP_PROLONG_ADD, op.level, op.omega, _FIELD[op.src or "sol"])
                elif k == "prolong":
                    add(MG_OP_PROLONG, op.level, 1.0, _FIELD[op.src or "sol"])
                elif k == "krylov":
                    if op.method not in MG_KRYLOV_METHOD:
                        raise EngineError(f"Krylov subspace method '{op.method}' is not executable")
                    add(MG_OP_KRYLOV, op.level, float(op.iterations), MG_KRYLOV_METHOD[op.method])
                elif k == "cgs":
                    if op.level == self.min_level:
                        add(MG_OP_CGS, op.level)
                    else:
                        inner = op.inner if op.inner else (default_cycle(op.level) if default_cycle else None)
                        if inner is None:
                            raise EngineError(f"coarse-grid solver on level {op.level} above the workspace's "
                                              f"min level {self.min_level} needs an inner cycle")
                        add(MG_OP_ZERO, op.level)
                        walk(lowering.fuse(inner))
                else:
                    raise EngineError(f"op kind '{k}' is not executable")

        walk(ops)
        return (MgOp * len(flat))(*flat)

    # -- data ------------------------------------------------------------------------------------
    def shape(self, level):
        """Host array shape of a field on `level` (slab: [owned planes + 2][n][n])."""
        n = 2 ** level + 1
        if self.slab:
            zb, ze = self.slab[level]
            return (ze - zb + 3, n, n)
        return (n,) * self.dimension

    def _dense(self, level, array):
        a = np.ascontiguousarray(array, dtype=self.np_dtype)
        if a.shape != self.shape(level):
            raise EngineError(f"array shape {a.shape} does not match level {level} ({self.shape(level)})")
        return a

    def field_view(self, kind: str, level: int, field: int):
        """Zero-copy torch view [nz][n][pitch] of a device field array (current solution buffer for
        "sol"): planes are contiguous, so a halo exchange sends `view[k]` as it is."""
        import torch
        ptr, n, pitch, nz, z0 = C.c_void_p(), C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.mg_field_pointer(self._h, _FIELD[kind], level, field, C.byref(ptr), C.byref(n),
                                              C.byref(pitch), C.byref(nz), C.byref(z0)))
        shape = (nz.value, n.value, pitch.value) if self.dimension == 3 else (n.value, pitch.value)

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": shape, "typestr": "<c16" if self.dtype == "c128" else "<f8",
                                        "data": (ptr.value, False), "version": 2}
        return torch.as_tensor(raw, device=torch.device("cuda", self.device))

    def upload(self, kind: str, level: int, field: int, array):
        a = self._dense(level, array)
        self._check(self.lib.mg_upload(self._h, _FIELD[kind], level, field, a.ctypes.data_as(C.c_void_p)))

    def upload_device(self, kind: str, level: int, field: int, device_ptr: int):
        self._check(self.lib.mg_upload_device(self._h, _FIELD[kind], level, field, C.c_void_p(device_ptr)))

    def download(self, kind: str, level: int, field: int) -> np.ndarray:
        out = np.empty(self.shape(level), dtype=self.np_dtype)
        self._check(self.lib.mg_download(self._h, _FIELD[kind], level, field, out.ctypes.data_as(C.c_void_p)))
        return out

    def save_initial_guess(self):
        self._check(self.lib.mg_save_initial_guess(self._h))

    def restore_initial_guess(self):
        self._check(self.lib.mg_restore_initial_guess(self._h))

    # -- execution ---------------------------------------------------------------------------------
    def exec_ops(self, ops, default_cycle=None, fused=False):
        enc = self.encode(lowering.fuse(ops) if fused else ops, default_cycle)
        self._check(self.lib.mg_exec_ops(self._h, enc, len(enc)))

    def enqueue_ops(self, ops, default_cycle=None, fused=False):
        """Issue the ops on the workspace stream without waiting (see set_stream)."""
        enc = self.encode(lowering.fuse(ops) if fused else ops, default_cycle)
        self._check(self.lib.mg_enqueue_ops(self._h, enc, len(enc)))

    def set_stream(self, cuda_stream: int):
        """Run on the caller's CUDA stream (e.g. torch.cuda.current_stream().cuda_stream); 0 = own stream."""
        self._check(self.lib.mg_set_stream(self._h, C.c_void_p(cuda_stream or None)))

    def build_cycle(self, ops, default_cycle=None, fused=True) -> int:
        enc = self.encode(lowering.fuse(ops) if fused else ops, default_cycle)
        cycle = C.c_int()
        self._check(self.lib.mg_build_cycle(self._h, enc, len(enc), C.byref(cycle)))
        return cycle.value

    def capture_begin(self):
        self._check(self.lib.mg_capture_begin(self._h))

    def capture_end(self) -> int:
        handle = C.c_int()
        self._check(self.lib.mg_capture_end(self._h, C.byref(handle)))
        return handle.value

    def graph_launch(self, graph: int, n: int = 1) -> bool:
        """-> False when the solution buffers are not in the state the graph was captured in."""
        rc = self.lib.mg_graph_launch(self._h, graph, n)
        if rc == -3:                       # MG_ERR_STATE
            return False
        self._check(rc)
        return True

    def run_cycles(self, cycle: int, n: int) -> float:
        ms = C.c_float()
        self._check(self.lib.mg_run_cycles(self._h, cycle, n, C.byref(ms)))
        return ms.value

    def residual_norm(self, level=None) -> float:
        out = C.c_double()
        self._check(self.lib.mg_residual_norm(self._h, self.max_level if level is None else level, C.byref(out)))
        return out.value

    def solve(self, cycle: int, tol: float, max_its: int):
        """-> (residual history, converged, ms)."""
        hist = (C.c_double * (max_its + 1))()
        n_its, conv, ms = C.c_int(), C.c_int(), C.c_float()
        self._check(self.lib.mg_solve(self._h, cycle, tol, max_its, hist, C.byref(n_its), C.byref(conv), C.byref(ms)))
        return list(hist[: n_its.value + 1]), bool(conv.value), ms.value

    def solve_bicgstab(self, cycle: int, tol: float, max_its: int):
        """Preconditioned BiCGStab with the cycle as preconditioner -> (history, converged, ms)."""
        hist = (C.c_double * (max_its + 1))()
        n_its, conv, ms = C.c_int(), C.c_int(), C.c_float()
        self._check(self.lib.mg_solve_bicgstab(self._h, cycle, tol, max_its, hist, C.byref(n_its), C.byref(conv),
                                               C.byref(ms)))
        return list(hist[: n_its.value + 1]), bool(conv.value), ms.value

    def synchronize(self):
        self._check(self.lib.mg_synchronize(self._h))

    def set_time_limit(self, seconds):
        """Wall-clock cap of one solve; an expired solve raises EngineError(code=MG_ERR_TIMEOUT)."""
        self._check(self.lib.mg_set_time_limit(self._h, float(seconds or 0.0)))

    def launch_count(self) -> int:
        out = C.c_int64()
        self._check(self.lib.mg_launch_count(self._h, C.byref(out)))
        return out.value

    # -- measurement -------------------------------------------------------------------------------
    def time_ops(self, ops, reps=1, default_cycle=None, fused=False) -> float:
        """Device time (ms) of `reps` back-to-back executions of the op list."""
        enc = self.encode(lowering.fuse(ops) if fused else ops, default_cycle)
        ms = C.c_float()
        self._check(self.lib.mg_time_ops(self._h, enc, len(enc), reps, C.byref(ms)))
        return ms.value

    def profile_ops(self, ops, default_cycle=None, fused=True):
        """-> [(kind, level, smoother handle, ms)] for one pass over the (flattened, fused) op list."""
        arr = self.encode(lowering.fuse(ops) if fused else ops, default_cycle)
        ms = (C.c_float * len(arr))()
        self._check(self.lib.mg_profile_ops(self._h, arr, len(arr), ms))
        return [(int(o.kind), int(o.level), int(o.smoother), float(t)) for o, t in zip(arr, ms)]

    def mark(self, which: int):
        self._check(self.lib.mg_mark(self._h, which))

    def elapsed(self) -> float:
        ms = C.c_float()
        self._check(self.lib.mg_elapsed(self._h, C.byref(ms)))
        return ms.value

    def flush_l2(self, nbytes=256 << 20):
        self._check(self.lib.mg_flush_l2(self._h, nbytes))

    # -- domain decomposition with peer-to-peer halo exchange (mg_dd.cuh, mg_engine_dd.inc) ----------------
    def dd_init(self, rank: int, world: int, coarse: "Workspace"):
        self._check(self.lib.mg_dd_init(self._h, rank, world, coarse._h))

    def dd_describe(self, level: int, ipc: bool):
        """-> (descriptor tuple of 8 ints, 64-byte CUDA IPC handle or None); level < 0: the comm block."""
        desc = (C.c_uint64 * 8)()
        handle = C.create_string_buffer(64) if ipc else None
        self._check(self.lib.mg_dd_describe(self._h, level, desc, handle))
        return tuple(int(v) for v in desc), (handle.raw if ipc else None)

    def dd_attach(self, peer: int, level: int, desc, handle):
        d = (C.c_uint64 * 8)(*desc)
        buf = C.create_string_buffer(handle, 64) if handle is not None else None
        self._check(self.lib.mg_dd_attach(self._h, peer, level, d, buf))

    def dd_exchange(self, kind: str, level: int, planes: int):
        self._check(self.lib.mg_dd_exchange(self._h, _FIELD[kind], level, planes))

    def dd_build_cycle(self, encoded) -> int:
        """`encoded`: MgOp array of the unfused cycle (domain.encode_for_slabs)."""
        cycle = C.c_int()
        self._check(self.lib.mg_dd_build_cycle(self._h, encoded, len(encoded), C.byref(cycle)))
        return cycle.value

    def dd_prepare(self, cycle: int):
        self._check(self.lib.mg_dd_prepare(self._h, cycle))

    def dd_launch(self, cycle: int, n: int = 1):
        self._check(self.lib.mg_dd_launch(self._h, cycle, n))

    def dd_norm(self):
        self._check(self.lib.mg_dd_norm(self._h))

    def dd_read_norm(self) -> float:
        out = C.c_double()
        self._check(self.lib.mg_dd_read_norm(self._h, C.byref(out)))
        return out.value

    # -- nonlinear FAS stages (mg_fas.cuh; FAS_2D_Basic.exa4) ------------------------------------
    def fas_smooth(self, level, omega, gamma, sweeps=1, newton_steps=1, colored=False):
        self._check(self.lib.mg_fas_smooth(self._h, level, omega, gamma, sweeps, newton_steps, 1 if colored else 0))

    def fas_residual(self, level, gamma):
        self._check(self.lib.mg_fas_residual(self._h, level, gamma))

    def fas_restrict(self, coarse_level, gamma):
        self._check(self.lib.mg_fas_restrict(self._h, coarse_level, gamma))

    def fas_subtract_approximation(self, level):
        self._check(self.lib.mg_fas_subtract_approximation(self._h, level))

    def fas_residual_norm(self, level, gamma) -> float:
        out = C.c_double()
        self._check(self.lib.mg_fas_residual_norm(self._h, level, gamma, C.byref(out)))
        return out.value

    def set_fast_kernels(self, enabled: bool, tail=None):
        """Specialised kernels on/off; `tail` (default: same as `enabled`) selects the single-CTA
        interpreter for the small levels."""
        tail = enabled if tail is None else tail
        self._check(self.lib.mg_set_fast_kernels(self._h, (1 if enabled else 0) | (2 if tail else 0)))

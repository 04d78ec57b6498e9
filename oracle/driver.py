"""ORACLE driver (test infrastructure): problem spec + lowered program -> oracle run.

Mirrors what the reference's generated `Application` does around the solver
(InitFields, apply bc, Solve; e.g. Helmholtz/2D_FD_Helmholtz_fromL3.exa4:1-22)."""
from __future__ import annotations

import numpy as np

from evostencils_b200 import lowering
from . import mg_oracle as mo


def make_hierarchy(problem, min_level, max_level, operators=None, params=None, engine_min_level=None):
    """`engine_min_level` < min_level adds the levels the default cycle below the
    grammar's coarse-grid solver needs."""
    lo = min_level if engine_min_level is None else engine_min_level
    p = dict(problem.parameters)
    p.update(params or {})
    dtype = np.complex128 if problem.dtype == "c128" else np.float64
    bc = mo.BoundaryCondition(problem.boundary, p.get("k", 0.0))
    h = mo.Hierarchy(problem.dimension, problem.n_fields, lo, max_level, dtype, bc)
    for level in range(lo, max_level + 1):
        table = lowering.operator_table(problem.system_operator(level, 0, p))
        h.operators[level] = table
    vc = problem.variable_coefficient_spec(p)
    if vc:
        # the IR carries the stencil at one position; the solver evaluates it at every node
        for level in range(lo, max_level + 1):
            h.operators[level] = mo.variable_poisson_table(problem.dimension, level, vc["kappa"])
            h.variable_levels.add(level)
        return h
    if operators:
        for level, table in operators.items():
            assert h.operators[level] == table, f"program operator on level {level} differs from the problem's"
    return h


def settings_for(problem, engine_min_level=None, params=None, max_iterations=None):
    s = problem.solver
    st = mo.SolverSettings(target_reduction=s.target_reduction,
                           max_iterations=s.max_iterations if max_iterations is None else max_iterations,
                           cgs=s.cgs, cgs_max_iterations=s.cgs_max_iterations,
                           cgs_target_reduction=s.cgs_target_reduction)
    if engine_min_level is not None:
        st.default_cycle = lambda level: lowering.default_cycle_ops(problem, level, engine_min_level, params)
    return st


def load_problem_fields(h, problem, params=None, rhs=None, u0=None):
    us, fs = problem.initial_fields(h.max_level, params)
    for i in range(h.n_fields):
        h.sol[h.max_level][i][...] = us[i] if u0 is None else u0[i]
        h.rhs[h.max_level][i][...] = fs[i] if rhs is None else rhs[i]
    h.apply_bc(h.max_level, h.sol[h.max_level])


def solve(problem, program, params=None, rhs=None, u0=None, engine_min_level=None, max_iterations=None,
          infinity=1e100):
    """-> (hierarchy, residual history, convergence factor, iterations)."""
    h = make_hierarchy(problem, program.min_level, program.max_level, program.operators, params, engine_min_level)
    st = settings_for(problem, engine_min_level, params, max_iterations)
    load_problem_fields(h, problem, params, rhs, u0)
    if problem.outer == "bicgstab":
        p = dict(problem.parameters)
        p.update(params or {})
        from evostencils_b200 import stencils
        a_table = [[lowering._canon(stencils.helmholtz_2d(problem.spacing(program.max_level), p["k"], 1.0))]]
        rhs_field = h.rhs[h.max_level][0].copy()
        hist, rho, n = mo.helmholtz_bicgstab_solve(h, program.ops, st, a_table, rhs_field, infinity,
                                                   problem.solver_iteration_limit)
    else:
        hist, rho, n = mo.mg_solve(h, program.ops, st, infinity, problem.solver_iteration_limit)
    return h, hist, rho, n

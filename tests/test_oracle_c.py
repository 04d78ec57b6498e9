"""The C/OpenMP oracle port (CPU baseline) reproduces the numpy oracle bit for bit."""
import os
import subprocess

import numpy as np
import pytest

from evostencils_b200 import engine, lowering, problems
from evostencils_b200.grammar import Grammar, v_cycle_string
from oracle import driver

ORACLE_DIR = os.path.join(os.path.dirname(__file__), "..", "oracle")
LIB = os.path.join(ORACLE_DIR, "_build", "libmgoracle.so")


@pytest.fixture(scope="module")
def clib():
    from oracle import cport
    return cport.load()


def c_solve(clib, problem, program, engine_min_level=None):
    lo = program.min_level if engine_min_level is None else engine_min_level
    ws = engine.Workspace(problem.dimension, lo, program.max_level, 1, "f64", 0, lib=clib)
    for level in range(lo, program.max_level + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver("CG", problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    us, fs = problem.initial_fields(program.max_level)
    ws.upload("sol", program.max_level, 0, us[0])
    ws.upload("rhs", program.max_level, 0, fs[0])
    default = None
    if engine_min_level is not None:
        default = lambda level: lowering.default_cycle_ops(problem, level, engine_min_level)
    cycle = ws.build_cycle(program.ops, default, fused=False)
    hist, conv, ms = ws.solve(cycle, problem.solver.target_reduction, problem.solver.max_iterations)
    u = ws.download("sol", program.max_level, 0)
    ws.close()
    return hist, conv, u


CASES = [
    (problems.POISSON_2D, 3, 7, lambda d: v_cycle_string(d, 2, 1, 21, "red_black"), None),
    (problems.POISSON_2D, 3, 6, lambda d: v_cycle_string(d, 2, 1, 14, "single"), None),
    (problems.POISSON_3D, 2, 5, lambda d: v_cycle_string(d, 2, 1, 23, "red_black"), None),
    (problems.POISSON_2D, 4, 7, lambda d: v_cycle_string(d, 1, 1, 21, "red_black"), 2),
]


@pytest.mark.parametrize("problem,lo,hi,make,engine_lo", CASES)
def test_c_port_matches_numpy_oracle(clib, problem, lo, hi, make, engine_lo):
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    program = lowering.lower(grammar.compile(make(hi - lo))[0], lo, hi)
    hist, conv, u = c_solve(clib, problem, program, engine_lo)
    h, want, rho, n = driver.solve(problem, program, engine_min_level=engine_lo)
    assert len(hist) == len(want)
    assert np.allclose(hist, want, rtol=1e-12, atol=0)
    assert np.array_equal(u, h.sol[hi][0])


def test_c_port_block_smoother(clib):
    import json
    g = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "lowering_golden.json")))["poisson2d_block22"]
    problem = problems.get_problem(g["problem"])
    grammar = Grammar(problem, g["min_level"], g["max_level"], maximum_local_system_size=4)
    program = lowering.lower(grammar.compile(g["grammar"])[0], g["min_level"], g["max_level"])
    hist, conv, u = c_solve(clib, problem, program)
    h, want, rho, n = driver.solve(problem, program)
    assert len(hist) == len(want)
    assert np.allclose(hist, want, rtol=1e-10, atol=0)
    assert np.allclose(u, h.sol[g["max_level"]][0], rtol=1e-11, atol=1e-13)

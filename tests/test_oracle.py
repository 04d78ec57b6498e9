"""Self-checks of the numpy oracle in lieu of golden vectors (the reference holds none for this
path): dense-matrix identities on small grids, the transfer-operator identity,
O(h^2) discretisation error against the analytic solution of config 1, textbook V-cycle factors."""
import itertools
import math

import numpy as np
import pytest
import scipy.sparse as sp

from evostencils_b200 import ir, lowering, problems
from evostencils_b200.grammar import Grammar, v_cycle_string
from oracle import driver, mg_oracle as mo


def dense_operator(table_entries, n, d):
    """Sparse matrix of one scalar stencil on the interior nodes of an n^d grid (zero Dirichlet)."""
    m = n - 2
    idx = np.arange(m ** d).reshape((m,) * d)
    rows, cols, vals = [], [], []
    for off, val in table_entries:                       # off = (x, y[, z]); array axes are reversed
        shift = tuple(reversed(off))
        src = [slice(max(0, -s), m - max(0, s)) for s in shift]
        dst = [slice(max(0, s), m - max(0, -s)) for s in shift]
        # row node p couples to node p + off
        rows.append(idx[tuple(src)].ravel())
        cols.append(idx[tuple(dst)].ravel())
        vals.append(np.full(rows[-1].shape, val))
    return sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(m ** d,) * 2)


def random_hierarchy(problem, lo, hi, seed=0):
    h = driver.make_hierarchy(problem, lo, hi)
    rng = np.random.default_rng(seed)
    for level in range(lo, hi + 1):
        for f in range(problem.n_fields):
            interior = tuple([slice(1, -1)] * problem.dimension)
            h.sol[level][f][interior] = rng.standard_normal(h.sol[level][f][interior].shape)
            h.rhs[level][f][interior] = rng.standard_normal(h.rhs[level][f][interior].shape)
    return h


@pytest.mark.parametrize("problem,level", [(problems.POISSON_2D, 4), (problems.POISSON_3D, 3)])
def test_residual_and_jacobi_match_dense_algebra(problem, level):
    d = problem.dimension
    n = 2 ** level + 1
    h = random_hierarchy(problem, level - 1, level)
    A = dense_operator(h.operators[level][0][0], n, d)
    interior = tuple([slice(1, -1)] * d)
    u = h.sol[level][0][interior].ravel().copy()
    f = h.rhs[level][0][interior].ravel().copy()
    mo.op_residual(h, level)
    assert np.allclose(h.res[level][0][interior].ravel(), f - A @ u, rtol=1e-12, atol=1e-9)
    op = problem.system_operator(level, 0)
    omega = 0.8
    stmts = lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, level, omega, ir.Single)
    mo.run_ops(h, stmts, driver.settings_for(problem))
    want = u + omega * (f - A @ u) / A.diagonal()
    assert np.allclose(h.sol[level][0][interior].ravel(), want, rtol=1e-12, atol=1e-12)


def test_red_black_equals_two_colour_product_form():
    """RB-GS = (I - w D^-1_B A)(I - w D^-1_R A) applied to the error (convergence.py:73-106 uses
    the same identity): colour 0 = (sum i_d) even first."""
    problem, level = problems.POISSON_2D, 4
    n = 2 ** level + 1
    h = random_hierarchy(problem, level - 1, level, seed=3)
    A = dense_operator(h.operators[level][0][0], n, 2)
    interior = (slice(1, -1), slice(1, -1))
    u = h.sol[level][0][interior].ravel().copy()
    f = h.rhs[level][0][interior].ravel().copy()
    yy, xx = np.indices((n - 2, n - 2)) + 1
    colour = ((xx + yy) % 2).ravel()
    omega = 1.15
    op = problem.system_operator(level, 0)
    stmts = lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, level, omega, ir.RedBlack)
    mo.run_ops(h, stmts, driver.settings_for(problem))
    for c in (0, 1):
        r = f - A @ u
        u = np.where(colour == c, u + omega * r / A.diagonal(), u)
    assert np.allclose(h.sol[level][0][interior].ravel(), u, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("problem,level", [(problems.POISSON_2D, 4), (problems.POISSON_3D, 3)])
def test_transfer_identity(problem, level):
    """<P v, w>_fine = 2^d <v, R w>_coarse for full weighting / multilinear interpolation."""
    d = problem.dimension
    h = random_hierarchy(problem, level - 1, level, seed=5)
    interior = tuple([slice(1, -1)] * d)
    v = h.sol[level - 1][0].copy()
    w = h.res[level][0]
    w[interior] = np.random.default_rng(7).standard_normal(w[interior].shape)
    h.sol[level][0][...] = 0
    mo.run_ops(h, [lowering.Op("prolong_add", level, omega=1.0, src="sol", dst="sol")], driver.settings_for(problem))
    lhs = np.sum(h.sol[level][0] * w)
    mo.run_ops(h, [lowering.Op("restrict", level - 1, src="res", dst="rhs")], driver.settings_for(problem))
    rhs = 2 ** d * np.sum(v * h.rhs[level - 1][0])
    assert abs(lhs - rhs) <= 1e-12 * abs(lhs)


def test_discretisation_error_is_second_order():
    """Config 1's PDE has the analytic solution u = cos(pi x) - sin(2 pi y)
    (Poisson/2D_FD_Poisson_fromL2.exa2:5-7): the converged discrete solution is O(h^2) away."""
    problem = problems.POISSON_2D
    errors = []
    for hi in (4, 5, 6):
        lo = 2
        grammar = Grammar(problem, lo, hi)
        expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))
        program = lowering.lower(expression, lo, hi)
        h, hist, rho, n = driver.solve(problem, program)
        assert n < 30 and hist[-1] < 1e-12 * hist[0]
        x = np.linspace(0.0, 1.0, 2 ** hi + 1)
        exact = np.cos(np.pi * x)[None, :] - np.sin(2 * np.pi * x)[:, None]
        errors.append(np.max(np.abs(h.sol[hi][0] - exact)))
    assert 3.5 < errors[0] / errors[1] < 4.5
    assert 3.5 < errors[1] / errors[2] < 4.5


def test_v_cycle_factors_are_textbook():
    """V(2,1) RB-GS on 2D Poisson with FW / bilinear transfer: ~0.03-0.12 per cycle; V(2,2)
    0.8-Jacobi: ~0.1-0.25 (sanity bands)."""
    problem, lo, hi = problems.POISSON_2D, 3, 6
    grammar = Grammar(problem, lo, hi)
    for string, band in ((v_cycle_string(hi - lo, 2, 1, 21, "red_black"), (0.02, 0.12)),
                         (v_cycle_string(hi - lo, 2, 2, 14, "single"), (0.08, 0.30))):
        expression, _ = grammar.compile(string)
        _, hist, rho, n = driver.solve(problem, lowering.lower(expression, lo, hi))
        assert band[0] < rho < band[1], rho
        assert n == len(hist) - 1


def test_summarise_conventions():
    """parse_output conventions (exastencils.py:539-584)."""
    from evostencils_b200.generator import summarise_history
    hist = [1.0, 0.1, 0.01, 0.001]
    rho, n = summarise_history(hist, 3, True)
    assert abs(rho - 0.1) < 1e-15 and n == 3.0
    rho, n = summarise_history(hist, 3, False)
    assert n == 1e100
    rho, n = summarise_history([1.0, float("nan")], 1, False)        # no finite factor at all -> infinity (:575-577)
    assert rho == 1e100 and n == 1e100
    rho, n = summarise_history([1.0, 0.5, float("nan")], 2, False)   # NaN factor counts as sqrt(infinity) (:549-551)
    assert rho == pytest.approx(math.sqrt(0.5 * 1e50)) and n == 1e100

"""Dense-matrix self-checks of the numpy oracle for the two configurations that are not scalar
Poisson: the 2-field elasticity block operator (config 5) and the complex Helmholtz operator with
slaved Robin columns (config 4).  tests/test_oracle.py holds the same identities for Poisson.

Every check builds the operator as an explicit matrix from the coefficients of the `.exa2/.exa3`
files (not from the oracle's stencil loops), applies textbook linear algebra, and compares with the
oracle's op -- so oracle and kernels cannot be wrong together on these paths without it showing."""
import numpy as np
import pytest

from evostencils_b200 import ir, lowering, problems
from oracle import driver, mg_oracle as mo


def interior_index(n):
    m = n - 2
    return np.arange(m * m).reshape(m, m)          # [y-1][x-1]


def assemble_block(table, n):
    """Dense F*m^2 x F*m^2 matrix of an F x F block of stencils on the interior of an n^2 grid,
    zero Dirichlet rim; unknown order field-major, then row-major (y, x)."""
    F, m = len(table), n - 2
    idx = interior_index(n)
    A = np.zeros((F * m * m, F * m * m), dtype=complex)
    for i in range(F):
        for j in range(F):
            for (ox, oy), val in table[i][j]:
                for y in range(m):
                    for x in range(m):
                        xx, yy = x + ox, y + oy
                        if 0 <= xx < m and 0 <= yy < m:
                            A[i * m * m + idx[y, x], j * m * m + idx[yy, xx]] += val
    return A


def flat(fields):
    return np.concatenate([f[1:-1, 1:-1].ravel() for f in fields])


def random_fields(h, level, rng, rim_zero=True):
    for f in range(h.n_fields):
        for arr in (h.sol[level][f], h.rhs[level][f]):
            v = rng.standard_normal(arr.shape)
            if np.iscomplexobj(arr):
                v = v + 1j * rng.standard_normal(arr.shape)
            arr[...] = v
            if rim_zero:
                arr[0, :] = arr[-1, :] = 0
                arr[:, 0] = arr[:, -1] = 0


# ---- elasticity: (lam+mu)(dxx u + dxy v) + lam Laplace u = f_u, same for v (LinearElasticity...exa2:45-50)

def elasticity_matrix_from_exa2(level):
    """The block matrix written down from the `.exa2` (stencils :19-43, equations :45-50), not from
    problems.py: central second differences, dxy = the 4-point corner stencil 1/(4 h^2)."""
    lam, mu = 195.0, 130.0
    n = 2 ** level + 1
    h2 = (1.0 / 2 ** level) ** 2
    dxx = [((-1, 0), 1 / h2), ((0, 0), -2 / h2), ((1, 0), 1 / h2)]
    dyy = [((0, -1), 1 / h2), ((0, 0), -2 / h2), ((0, 1), 1 / h2)]
    lap = [((0, 0), -2 / h2 - 2 / h2), ((-1, 0), 1 / h2), ((1, 0), 1 / h2), ((0, -1), 1 / h2), ((0, 1), 1 / h2)]
    dxy = [((-1, 1), -1 / (4 * h2)), ((1, 1), 1 / (4 * h2)), ((-1, -1), 1 / (4 * h2)), ((1, -1), -1 / (4 * h2))]

    def combo(*terms):
        out = {}
        for factor, sten in terms:
            for off, v in sten:
                out[off] = out.get(off, 0.0) + factor * v
        return list(out.items())
    table = [[combo((lam + mu, dxx), (lam, lap)), combo((lam + mu, dxy))],
             [combo((lam + mu, dxy)), combo((lam + mu, dyy), (lam, lap))]]
    return assemble_block(table, n).real, n


def test_elasticity_operator_values_follow_the_exa2():
    """problems.py's block operator equals the matrix written down from the `.exa2` coefficients."""
    level = 3
    want, n = elasticity_matrix_from_exa2(level)
    problem = problems.LINEAR_ELASTICITY_2D
    table = lowering.operator_table(problem.system_operator(level, 0))
    got = assemble_block(table, n).real
    assert np.allclose(got, want, rtol=1e-13, atol=1e-9)


@pytest.mark.parametrize("kind", ["collective", "decoupled"])
@pytest.mark.parametrize("partitioning", ["single", "red_black"])
def test_elasticity_point_smoothers_match_dense_algebra(kind, partitioning):
    problem, level = problems.LINEAR_ELASTICITY_2D, 3
    A, n = elasticity_matrix_from_exa2(level)
    m = n - 2
    h = driver.make_hierarchy(problem, level - 1, level)
    random_fields(h, level, np.random.default_rng(11))
    u, f = flat(h.sol[level]), flat(h.rhs[level])
    mo.op_residual(h, level)
    assert np.allclose(flat(h.res[level]), f - A @ u, rtol=1e-12, atol=1e-7)
    omega = 0.9
    op = problem.system_operator(level, 0)
    marker = ir.ElementwiseDiagonal(op) if kind == "collective" else ir.Diagonal(op)
    part = ir.Single if partitioning == "single" else ir.RedBlack
    mo.run_ops(h, lowering.local_solve_statements(marker, op, level, omega, part), driver.settings_for(problem))
    # dense: point-block diagonal D (2x2 per node, or its diagonal), colour by colour
    D = np.zeros_like(A)
    for p in range(m * m):
        ids = [p, m * m + p]
        if kind == "collective":
            D[np.ix_(ids, ids)] = A[np.ix_(ids, ids)]
        else:
            for i in ids:
                D[i, i] = A[i, i]
    Dinv = np.linalg.inv(D)
    yy, xx = np.indices((m, m)) + 1
    colour = np.tile(((xx + yy) % 2).ravel(), 2)
    if partitioning == "single":
        if kind == "collective":
            u = u + omega * Dinv @ (f - A @ u)
        else:                       # decoupled: one statement per field, each on the current values
            for fld in range(2):
                sel = np.zeros(2 * m * m, bool)
                sel[fld * m * m:(fld + 1) * m * m] = True
                u = np.where(sel, u + omega * Dinv @ (f - A @ u), u)
    else:
        if kind == "collective":
            for c in (0, 1):
                u = np.where(colour == c, u + omega * Dinv @ (f - A @ u), u)
        else:
            for fld in range(2):
                sel = np.zeros(2 * m * m, bool)
                sel[fld * m * m:(fld + 1) * m * m] = True
                for c in (0, 1):
                    u = np.where(sel & (colour == c), u + omega * Dinv @ (f - A @ u), u)
    assert np.allclose(flat(h.sol[level]), u, rtol=1e-11, atol=1e-11)


def test_elasticity_block_jacobi_matches_dense_algebra():
    """collective_block_jacobi with block shape ((2,1),(2,1)): tiles of two x-adjacent
    nodes times two fields, anchored at node 0, the local 4x4 (or clipped) system solved exactly."""
    problem, level = problems.LINEAR_ELASTICITY_2D, 3
    A, n = elasticity_matrix_from_exa2(level)
    m = n - 2
    idx = interior_index(n)
    h = driver.make_hierarchy(problem, level - 1, level)
    random_fields(h, level, np.random.default_rng(12))
    u, f = flat(h.sol[level]), flat(h.rhs[level])
    omega = 0.7
    op = problem.system_operator(level, 0)
    stmts = lowering.local_solve_statements(ir.BlockDiagonalOperator(op, ((2, 1), (2, 1))), op, level, omega, ir.Single)
    mo.run_ops(h, stmts, driver.settings_for(problem))
    r = f - A @ u
    new = u.copy()
    for y in range(1, n - 1):
        for tx in range((n - 1) // 2 + 1):
            nodes = [x for x in (2 * tx, 2 * tx + 1) if 1 <= x <= n - 2]
            if not nodes:
                continue
            ids = [fld * m * m + idx[y - 1, x - 1] for fld in range(2) for x in nodes]
            new[ids] = u[ids] + omega * np.linalg.solve(A[np.ix_(ids, ids)], r[ids])
    assert np.allclose(flat(h.sol[level]), new, rtol=1e-11, atol=1e-11)


# ---- Helmholtz: M = -Laplace - (1 + 0.5 i) k^2, Robin columns slaved: u_0 = u_1 / (1 - i k h) -----------

def helmholtz_effective_matrix(level, k, shift):
    """Interior matrix with the boundary eliminated: rows y = 0, n-1 are Dirichlet 0, the columns
    x = 0, n-1 hold c times their interior neighbour (Helmholtz/...exa4:25-59), so the coupling of an
    interior node x = 1 to its left neighbour folds onto its own diagonal with the factor c."""
    n = 2 ** level + 1
    m = n - 2
    h = 1.0 / 2 ** level
    c = 1.0 / (1.0 - 1j * k * h)
    idx = interior_index(n)
    M = np.zeros((m * m, m * m), dtype=complex)
    centre = 4.0 / h ** 2 - shift * k ** 2
    off = -1.0 / h ** 2
    for y in range(m):
        for x in range(m):
            p = idx[y, x]
            M[p, p] += centre
            for dx, dy in ((-1, 0), (1, 0), (0, -1), (0, 1)):
                xx, yy = x + dx, y + dy
                if 0 <= xx < m and 0 <= yy < m:
                    M[p, idx[yy, xx]] += off
                elif 0 <= yy < m:                 # x-boundary: slaved to this very node
                    M[p, p] += off * c
    return M, n, c


def test_helmholtz_operator_and_jacobi_match_dense_algebra():
    problem, level, k = problems.HELMHOLTZ_2D, 4, 20.0
    M, n, c = helmholtz_effective_matrix(level, k, 1.0 + 0.5j)
    h = driver.make_hierarchy(problem, level - 1, level, params={"k": k})
    random_fields(h, level, np.random.default_rng(13))
    h.apply_bc(level, h.sol[level])
    # the slaved columns hold c times their neighbour, the Dirichlet rows zero
    assert np.allclose(h.sol[level][0][1:-1, 0], c * h.sol[level][0][1:-1, 1])
    assert np.all(h.sol[level][0][0, :] == 0)
    u, f = flat(h.sol[level]), flat(h.rhs[level])
    mo.op_residual(h, level)
    assert np.allclose(flat(h.res[level]), f - M @ u, rtol=1e-12, atol=1e-8)
    # damped Jacobi uses the stencil's centre coefficient (not the folded diagonal), then re-slaves
    omega = 0.6
    op = problem.system_operator(level, 0, {"k": k})
    a00 = 4.0 * 4 ** level - (1.0 + 0.5j) * k * k
    stmts = lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, level, omega, ir.Single)
    mo.run_ops(h, stmts, driver.settings_for(problem))
    want = u + omega * (f - M @ u) / a00
    assert np.allclose(flat(h.sol[level]), want, rtol=1e-12, atol=1e-12)
    assert np.allclose(h.sol[level][0][1:-1, -1], c * h.sol[level][0][1:-1, -2])


def test_helmholtz_coarse_bicgstab_solves_the_slaved_system():
    """The coarse solve (Helmholtz/...exa3:396-433, reduction 1e-6) returns x with
    |f - M_eff x| <= 1e-6 |f| for the matrix with the Robin columns eliminated."""
    problem, level, k = problems.HELMHOLTZ_2D, 3, 10.0
    M, n, c = helmholtz_effective_matrix(level, k, 1.0 + 0.5j)
    h = driver.make_hierarchy(problem, level, level + 1, params={"k": k})
    random_fields(h, level, np.random.default_rng(14))
    f = flat(h.rhs[level])
    its = mo.op_coarse_bicgstab(h, level, 1000, 1e-6)
    x = flat(h.sol[level])
    assert 0 < its < 200
    # the stopping rule uses the unconjugated "norm" sqrt(r.r); the Euclidean residual is of that order
    assert np.linalg.norm(f - M @ x) <= 5e-6 * np.linalg.norm(f)
    assert np.allclose(x, np.linalg.solve(M, f), rtol=0, atol=1e-4 * np.abs(x).max())


def test_explicit_complex_arithmetic_agrees_with_numpy():
    """mul / div / root spell out the textbook formulas (bit-reproducible across machines); they
    agree with numpy's complex arithmetic to rounding."""
    rng = np.random.default_rng(15)
    a = rng.standard_normal(50) + 1j * rng.standard_normal(50)
    b = rng.standard_normal(50) + 1j * rng.standard_normal(50)
    assert np.allclose(mo.mul(a, b), a * b, rtol=1e-14, atol=1e-15)
    assert np.allclose(mo.mul(a, 2.5), a * 2.5, rtol=1e-15)
    assert np.allclose(mo.scale(0.25, a), 0.25 * a, rtol=0, atol=0)
    for x, y in zip(a[:10], b[:10]):
        assert abs(mo.div(complex(x), complex(y)) - x / y) <= 1e-14 * abs(x / y)
        assert abs(mo.root(complex(x)) - np.sqrt(x)) <= 1e-14 * abs(np.sqrt(x))
        assert abs(mo.modulus(complex(x)) - abs(x)) <= 1e-15 * abs(x)
    assert mo.mul(3.0, 4.0) == 12.0 and mo.div(3.0, 4.0) == 0.75 and mo.root(4.0) == 2.0


def test_grid_dot_is_the_plain_interior_sum_reordered():
    rng = np.random.default_rng(16)
    for n in (9, 129, 300):
        a = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
        b = rng.standard_normal((n, n)) + 1j * rng.standard_normal((n, n))
        want = np.sum(a[1:-1, 1:-1] * b[1:-1, 1:-1])
        assert abs(mo.grid_dot(a, b) - want) <= 1e-12 * abs(want)

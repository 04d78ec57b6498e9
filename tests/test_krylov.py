"""Krylov subspace methods in the place of a smoother (ir/krylov_subspace.py:4-45, emitted at
code_generation/exastencils.py:744-768).

No golden listing exists for this construct: the reference's node cannot be constructed
(`KrylovSubspaceMethod.__init__` calls `Entity.__init__()` without its three arguments -> TypeError,
checked below when the reference is present), its grammar has no production for it and the solver functions
the emitted call refers to come out of an ExaStencils run (:1074-1101, "either fix or remove").  The
statement text is therefore checked against the emitter's source lines, the semantics are the engine's own
(oracle.mg_oracle.op_krylov) and the GPU is compared with the oracle."""
import numpy as np
import pytest

from evostencils_b200 import exaslang, ir, lowering, problems
from evostencils_b200.grammar import Grammar

from conftest import REFERENCE, have_reference


def krylov_string(depth, method, iterations, relax_index=21):
    """V-cycle: `iterations` steps of `method` as pre-smoother, one red-black GS sweep as post-smoother."""
    expr = "u_and_f"
    for d in range(depth):
        inner = f"residual_{d}({expr})" if d == 0 else expr          # coarsening hands over a cycle with its residual
        expr = f"residual_{d}({method}_{d}(its_{iterations}, {inner}))"
        if d < depth - 1:
            expr = f"coarsening_{d}(A_{d + 1}, zero_{d + 1}, R_{d}, {expr})"
    d = depth - 1
    expr = f"correct_with_coarse_grid_solver_{d}(18, P_{d + 1}, CGS_{d + 1}, R_{d}, {expr})"
    for d in range(depth - 1, -1, -1):
        expr = f"collective_jacobi_{d}({relax_index}, red_black, residual_{d}({expr}))"
        if d > 0:
            expr = f"update_with_coarse_grid_correction_{d - 1}(18, P_{d}, {expr})"
    return expr


CASES = {
    "poisson2d_cg": (problems.POISSON_2D, 2, 5, "conjugate_gradient", 4, 21),
    "poisson3d_cg": (problems.POISSON_3D, 2, 4, "conjugate_gradient", 2, 23),
    "elasticity_bicgstab": (problems.LINEAR_ELASTICITY_2D, 3, 5, "bicgstab", 2, 23),
    "helmholtz_bicgstab": (problems.HELMHOLTZ_2D, 3, 5, "bicgstab", 4, 10),
}


def build(name):
    problem, lo, hi, method, its, relax = CASES[name]
    grammar = Grammar(problem, lo, hi, krylov_methods=("conjugate_gradient", "bicgstab", "minres"),
                      krylov_iterations=(2, 4, 8), parameters={"k": 10.0} if problem.outer == "bicgstab" else None)
    expression, _ = grammar.compile(krylov_string(hi - lo, method, its, relax))
    return problem, lo, hi, method, its, expression, grammar


def test_reference_node_cannot_be_constructed():
    if not have_reference():
        pytest.skip("reference checkout not present")
    import sys
    sys.path.insert(0, REFERENCE)
    from evostencils.ir import krylov_subspace
    class _Operator:
        shape = (9, 9)
        grid = None

    with pytest.raises(TypeError):
        krylov_subspace.generate_conjugate_gradient(_Operator(), 4)


@pytest.mark.parametrize("name", sorted(CASES))
def test_lowering_and_exaslang(name):
    problem, lo, hi, method, its, expression, _ = build(name)
    program = lowering.lower(expression, lo, hi)
    krylov = [op for op in program.ops if getattr(op, "kind", "") == "krylov"]
    assert [op.level for op in krylov] == list(range(hi, lo, -1))
    node = {"conjugate_gradient": "ConjugateGradient", "bicgstab": "BiCGStab"}[method]
    assert all(op.method == node and op.iterations == its for op in krylov)
    # the statement in front of every Krylov op on a coarse level is the zero of its approximation
    for i, op in enumerate(program.ops):
        if getattr(op, "kind", "") == "krylov" and op.level < hi:
            assert program.ops[i - 1].kind == "zero" and program.ops[i - 1].level == op.level
    text = exaslang.emit_cycle_function(problem, expression, lo, hi, hi, "gen_mgCycle", dict(problem.parameters))
    lines = [line.strip() for line in text.splitlines()]
    for level in range(hi, lo, -1):
        call = f"{node}_{its}@{level}()"                       # exastencils.py:766
        assert call in lines
        at = lines.index(call)
        if level < hi:                                         # :761-764: gen_rhs_<field>@level = <rhs field>, one per field
            for k, field in enumerate(reversed(problem.field_names)):
                assert lines[at - 1 - k].startswith(f"gen_rhs_{field}@{level} = ")
        else:
            assert not lines[at - 1].startswith("gen_rhs_")
    assert lowering.estimate_cycle_bytes(program.ops, problem.dimension) > 0


def test_methods_without_an_executor_give_the_sentinel_path():
    problem, lo, hi = problems.POISSON_2D, 2, 4
    grammar = Grammar(problem, lo, hi, krylov_methods=("minres",), krylov_iterations=(2,))
    expression, _ = grammar.compile(krylov_string(hi - lo, "minres", 2))
    with pytest.raises(lowering.LoweringError):
        lowering.lower(expression, lo, hi)
    node = ir.generate_conjugate_residual(problem.system_operator(3, 0), 8)
    assert node.name == "ConjugateResidual" and node.number_of_iterations == 8 and node.generate_stencil() is None


def test_oracle_krylov_smoother_is_cg_started_from_the_iterate():
    """x += CG_n(A, rhs - A x) from zero equals n steps of textbook CG started at x."""
    from oracle import driver, mg_oracle as mo
    problem, level = problems.POISSON_2D, 4
    h = driver.make_hierarchy(problem, level, level)
    rng = np.random.default_rng(1)
    shape = (2 ** level + 1,) * 2
    u = rng.standard_normal(shape)
    f = rng.standard_normal(shape)
    h.sol[level][0][...] = u
    h.rhs[level][0][...] = f
    mo.op_krylov(h, level, "ConjugateGradient", 5, 1e-30)
    got = mo._interior(h.sol[level][0]).copy()

    def A(v):
        full = np.zeros(shape)
        mo._interior(full)[...] = v
        full[0, :], full[-1, :], full[:, 0], full[:, -1] = u[0, :], u[-1, :], u[:, 0], u[:, -1]
        return mo.apply_operator(h.operators[level], [full], 0)

    def A0(v):
        full = np.zeros(shape)
        mo._interior(full)[...] = v
        return mo.apply_operator(h.operators[level], [full], 0)

    x = mo._interior(u).copy()
    r = mo._interior(f) - A(x)
    p = r.copy()
    for _ in range(5):
        ap = A0(p)
        alpha = np.sum(r * r) / np.sum(p * ap)
        x = x + alpha * p
        r_new = r - alpha * ap
        p = r_new + (np.sum(r_new * r_new) / np.sum(r * r)) * p
        r = r_new
    assert np.max(np.abs(got - x)) <= 1e-11 * np.max(np.abs(x))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_solves_match_the_oracle(name):
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem, lo, hi, method, its, expression, _ = build(name)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.initialize_code_generation(lo, hi)
    program = lowering.lower(expression, lo, hi)
    params = dict(problem.parameters)
    if problem.outer == "bicgstab":
        params["k"] = 10.0
    # a Krylov method is not a stationary preconditioner: inside the outer BiCGStab of the Helmholtz problem the
    # iteration does not converge, so the first 12 outer iterations are compared there
    cap = 12 if problem.outer == "bicgstab" else None
    t, rho, n = gen.evaluate_program(program, params, evaluation_samples=1, max_iterations=cap)
    _, hist, rho_ref, n_ref = driver.solve(problem, lowering.lower(expression, lo, hi), params=params, max_iterations=cap)
    got, want = np.array(gen.last_residual_history), np.array(hist)
    assert n == n_ref and got.shape == want.shape, (n, n_ref)
    assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want) + 1e-14 * want[0]), (got, want)
    gen.close()


@pytest.mark.gpu
def test_gpu_krylov_stage_on_a_level_beyond_shared_memory():
    """129^2: the four CG vectors (4 x 133 KB) use the global scratch."""
    import test_gpu_parity as tp
    problem, lo, hi = problems.POISSON_2D, 6, 7
    ws, h = tp.make_pair(problem, lo, hi, seed=4)
    ops = [lowering.Op("krylov", hi, src="rhs", dst="sol", method="ConjugateGradient", iterations=6)]
    tp.run_both(ws, h, ops, problem)
    tp.compare(ws, h, "sol", hi, True, "CG smoother")
    ws.close()

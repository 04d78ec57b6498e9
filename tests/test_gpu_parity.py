"""GPU parity tests: every CUDA stage and whole solves against the numpy oracle, through the C ABI.

Bar (DESIGN.md "Numerical contract"): fields bit-identical to the oracle for point smoothers,
residual, restriction, prolongation and the coarse Krylov solves -- real and complex (the oracle
states complex products / quotients / roots in the engine's operand order); <= 1e-12 relative for
tile smoothers (LAPACK vs closed-form inverse is shared, summation identical -> in practice exact);
residual norms <= 1e-10 relative (north_star), iteration counts identical."""
import numpy as np
import pytest

from evostencils_b200 import ir, lowering, problems
from evostencils_b200.grammar import Grammar, v_cycle_string

pytestmark = pytest.mark.gpu

CASES = {
    "poisson2d": (problems.POISSON_2D, 3, 6),
    "poisson3d": (problems.POISSON_3D, 2, 4),
    "elasticity": (problems.LINEAR_ELASTICITY_2D, 3, 5),
    "helmholtz": (problems.HELMHOLTZ_2D, 3, 5),
}


def make_pair(problem, lo, hi, seed=0, use_fast=True):
    """Engine workspace + oracle hierarchy holding identical random data on every level."""
    from evostencils_b200.engine import Workspace
    from oracle import driver
    ws = Workspace(problem.dimension, lo, hi, problem.n_fields, problem.dtype, 0)
    h = driver.make_hierarchy(problem, lo, hi)
    p = dict(problem.parameters)
    for level in range(lo, hi + 1):
        ws.set_operator(level, h.operators[level])
    ws.set_boundary(problem.boundary, p.get("k", 0.0))
    ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    rng = np.random.default_rng(seed)
    for level in range(lo, hi + 1):
        shape = (2 ** level + 1,) * problem.dimension
        for f in range(problem.n_fields):
            def rnd():
                a = rng.standard_normal(shape)
                if problem.dtype == "c128":
                    a = a + 1j * rng.standard_normal(shape)
                return a.astype(h.dtype)
            u, r, s = rnd(), rnd(), rnd()
            interior = tuple([slice(1, -1)] * problem.dimension)
            if level < hi:               # error equations: zero Dirichlet rim
                rim = np.ones(shape, bool)
                rim[interior] = False
                u[rim] = 0
            h.sol[level][f][...] = u
            h.rhs[level][f][...] = r
            h.res[level][f][...] = s
        h.apply_bc(level, h.sol[level])
        for f in range(problem.n_fields):
            ws.upload("sol", level, f, h.sol[level][f])
            ws.upload("rhs", level, f, h.rhs[level][f])
            ws.upload("res", level, f, h.res[level][f])
    return ws, h


def compare(ws, h, kind, level, exact, what):
    for f in range(h.n_fields):
        got = ws.download(kind, level, f)
        want = h.field(kind, level)[f]
        if exact:
            bad = np.argwhere(got != want)
            assert bad.size == 0, f"{what}: {len(bad)} mismatching nodes, first {bad[:3].tolist()}, " \
                                  f"max abs diff {np.max(np.abs(got - want))}"
        else:
            scale = np.max(np.abs(want)) + 1e-300
            assert np.max(np.abs(got - want)) <= 1e-12 * scale, what


def run_both(ws, h, ops, problem, fused=False):
    from oracle import driver, mg_oracle
    ws.exec_ops(ops, fused=fused)
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))


def smoother_statements(problem, level, kind, partitioning, omega=0.8, block=None):
    op = problem.system_operator(level, 0)
    if kind == "collective":
        marker = ir.ElementwiseDiagonal(op)
    elif kind == "decoupled":
        marker = ir.Diagonal(op)
    else:
        marker = ir.BlockDiagonalOperator(op, block)
    return lowering.local_solve_statements(marker, op, level, omega, partitioning)


@pytest.mark.parametrize("name", sorted(CASES))
def test_residual_restrict_prolong(name):
    problem, lo, hi = CASES[name]
    exact = True
    ws, h = make_pair(problem, lo, hi, seed=1)
    run_both(ws, h, [lowering.Op("residual", hi, dst="res")], problem)
    compare(ws, h, "res", hi, exact, "residual")
    run_both(ws, h, [lowering.Op("restrict", hi - 1, src="res", dst="rhs")], problem)
    compare(ws, h, "rhs", hi - 1, exact, "restrict")
    run_both(ws, h, [lowering.Op("prolong_add", hi, omega=0.55, src="sol", dst="sol")], problem)
    compare(ws, h, "sol", hi, exact, "prolong_add")
    # fused residual+restriction equals the two separate stages
    ops = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    from oracle import driver, mg_oracle
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))
    ws.exec_ops(ops, fused=True)
    compare(ws, h, "rhs", hi - 1, exact, "residual_restrict")
    run_both(ws, h, [lowering.Op("zero", hi - 1, dst="sol")], problem)
    compare(ws, h, "sol", hi - 1, True, "zero")
    ws.close()


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("partitioning", ["single", "red_black"])
@pytest.mark.parametrize("kind", ["collective", "decoupled"])
def test_point_smoothers(name, partitioning, kind):
    problem, lo, hi = CASES[name]
    if kind == "decoupled" and problem.n_fields == 1:
        pytest.skip("decoupled == collective for one field")
    exact = True
    part = ir.Single if partitioning == "single" else ir.RedBlack
    ws, h = make_pair(problem, lo, hi, seed=2)
    for level in (hi, lo + 1):
        stmts = smoother_statements(problem, level, kind, part, omega=1.15)
        for sweep in range(3):          # odd number of Jacobi sweeps exercises the ping-pong state
            run_both(ws, h, stmts, problem)
        compare(ws, h, "sol", level, exact, f"{kind} {partitioning} level {level}")
    ws.close()


@pytest.mark.parametrize("name,block", [
    ("poisson2d", ((2, 2),)), ("poisson2d", ((1, 3),)), ("poisson2d", ((4, 1),)), ("poisson2d", ((3, 1),)),
    ("poisson3d", ((2, 1, 2),)), ("poisson3d", ((1, 1, 3),)),
    ("elasticity", ((2, 1), (2, 1))), ("elasticity", ((1, 2), (1, 2))),
    ("helmholtz", ((2, 2),)),
])
def test_tile_smoothers(name, block):
    problem, lo, hi = CASES[name]
    ws, h = make_pair(problem, lo, hi, seed=3)
    for level in (hi, lo):
        stmts = smoother_statements(problem, level, "block", ir.Single, omega=0.6, block=block)
        for sweep in range(2):
            run_both(ws, h, stmts, problem)
        compare(ws, h, "sol", level, False, f"block {block} level {level}")
    ws.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_coarse_solver_and_norm(name):
    problem, lo, hi = CASES[name]
    exact = True
    from oracle import mg_oracle
    ws, h = make_pair(problem, lo, hi, seed=4)
    for level in (hi, lo):
        got = ws.residual_norm(level)
        want = mg_oracle.residual_norm(h, level)
        assert abs(got - want) <= 1e-12 * want
    run_both(ws, h, [lowering.Op("cgs", lo, src="rhs", dst="sol")], problem)
    compare(ws, h, "sol", lo, exact, "coarse solve")
    ws.close()


def test_coarse_cg_on_the_config1_level():
    """CG on 33x33 (config 1: levels 5..9) -- hundreds of iterations, still bit-identical."""
    problem = problems.POISSON_2D
    ws, h = make_pair(problem, 5, 6, seed=5)
    run_both(ws, h, [lowering.Op("cgs", 5, src="rhs", dst="sol")], problem)
    compare(ws, h, "sol", 5, True, "coarse CG 33x33")
    ws.close()


SOLVES = [
    ("poisson2d_v21_rbgs", problems.POISSON_2D, 5, 9, lambda d: v_cycle_string(d, 2, 1, 21, "red_black")),
    ("poisson2d_v22_jacobi", problems.POISSON_2D, 4, 8, lambda d: v_cycle_string(d, 2, 2, 14, "single")),
    ("poisson2d_v11_jacobi_odd", problems.POISSON_2D, 3, 6, lambda d: v_cycle_string(d, 2, 1, 14, "single")),
    ("poisson3d_v21_rbgs", problems.POISSON_3D, 2, 6, lambda d: v_cycle_string(d, 2, 1, 23, "red_black")),
    ("elasticity_v21_rb", problems.LINEAR_ELASTICITY_2D, 4, 7, lambda d: v_cycle_string(d, 2, 1, 23, "red_black")),
    ("elasticity_v11_decoupled", problems.LINEAR_ELASTICITY_2D, 4, 6,
     lambda d: v_cycle_string(d, 1, 1, 14, "single", smoother="decoupled_jacobi")),
]


@pytest.mark.parametrize("name,problem,lo,hi,make", SOLVES, ids=[s[0] for s in SOLVES])
def test_solve_residual_history(name, problem, lo, hi, make):
    """north_star: identical iteration count, per-iteration residual norms <= 1e-10 relative."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    expression, _ = grammar.compile(make(hi - lo))
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.initialize_code_generation(lo, hi)
    t, rho, n = gen.generate_and_evaluate(expression, gen.generate_storage(lo, hi, gen.finest_grid), lo, hi, "",
                                          evaluation_samples=2)
    program = lowering.lower(expression, lo, hi)
    h, hist, rho_ref, n_ref = driver.solve(problem, program)
    got, want = np.array(gen.last_residual_history), np.array(hist)
    assert n == n_ref
    assert got.shape == want.shape
    assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want)), np.max(np.abs(got - want) / np.abs(want))
    assert abs(rho - rho_ref) <= 1e-10 * rho_ref
    assert t > 0
    # the converged iterate itself is bit-identical for real problems
    for f in range(problem.n_fields):
        u = gen._workspace.download("sol", hi, f)
        assert np.array_equal(u, h.sol[hi][f])
    gen.close()


def test_solve_with_golden_individuals():
    """The reference-emitted golden individuals (tests/golden) run and match the oracle."""
    import json, os
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "lowering_golden.json")))
    for case in ("survey_3_2", "poisson2d_block22", "poisson2d_wlike", "elasticity_block"):
        g = golden[case]
        problem = problems.get_problem(g["problem"])
        lo, hi = g["min_level"], g["max_level"]
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=g["maximum_local_system_size"])
        expression, _ = grammar.compile(g["grammar"])
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
        gen.initialize_code_generation(lo, hi)
        t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
        program = lowering.lower(expression, lo, hi)
        _, hist, rho_ref, n_ref = driver.solve(problem, program)
        got, want = np.array(gen.last_residual_history), np.array(hist)
        assert n == n_ref, case
        assert got.shape == want.shape, case
        assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want)), case
        gen.close()


def test_determinism():
    """Same input twice -> bitwise identical residual history (fixed reduction order)."""
    from evostencils_b200.generator import B200ProgramGenerator
    problem, lo, hi = problems.POISSON_2D, 4, 9
    grammar = Grammar(problem, lo, hi)
    expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))
    hists = []
    for _ in range(2):
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
        gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
        hists.append(list(gen.last_residual_history))
        gen.close()
    assert hists[0] == hists[1]


def test_inner_default_cycle_below_the_grammar():
    """Workspace deeper than the grammar: CGS on the grammar's coarsest level = one default cycle."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem, lo, hi, engine_lo = problems.POISSON_2D, 5, 8, 2
    grammar = Grammar(problem, lo, hi)
    expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, engine_min_level=engine_lo)
    t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
    program = lowering.lower(expression, lo, hi)
    _, hist, rho_ref, n_ref = driver.solve(problem, program, engine_min_level=engine_lo)
    got, want = np.array(gen.last_residual_history), np.array(hist)
    assert n == n_ref and got.shape == want.shape
    assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want))
    gen.close()


# ---- specialised (row-marching) 2D kernels: levels >= 8 -------------------------------------------

@pytest.mark.parametrize("lo,hi", [(7, 9), (9, 10)])
def test_fast_2d_stages_bit_identical(lo, hi):
    problem = problems.POISSON_2D
    ws, h = make_pair(problem, lo, hi, seed=7)
    run_both(ws, h, [lowering.Op("residual", hi, dst="res")], problem)
    compare(ws, h, "res", hi, True, "fast residual")
    from oracle import mg_oracle
    got, want = ws.residual_norm(hi), mg_oracle.residual_norm(h, hi)
    assert abs(got - want) <= 1e-12 * want
    for partitioning in (ir.Single, ir.RedBlack):
        for level in (hi, hi - 1):
            stmts = smoother_statements(problem, level, "collective", partitioning, omega=1.15)
            for sweep in range(3):
                run_both(ws, h, stmts, problem)
            compare(ws, h, "sol", level, True, f"fast smoother {partitioning.get_name()} level {level}")
    ops = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    from oracle import driver
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))
    ws.exec_ops(ops, fused=True)
    compare(ws, h, "rhs", hi - 1, True, "fast residual_restrict")
    run_both(ws, h, [lowering.Op("prolong_add", hi, omega=0.55, src="sol", dst="sol")], problem)
    compare(ws, h, "sol", hi, True, "fast prolong_add")
    ws.close()


def test_fast_and_generic_kernels_agree_on_a_solve():
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem, lo, hi = problems.POISSON_2D, 5, 10
    grammar = Grammar(problem, lo, hi)
    for string in (v_cycle_string(hi - lo, 2, 1, 21, "red_black"), v_cycle_string(hi - lo, 2, 1, 14, "single")):
        expression, _ = grammar.compile(string)
        hists = []
        for fast in (True, False):
            gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
            gen.initialize_code_generation(lo, hi)
            gen._workspace.set_fast_kernels(fast)
            gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
            hists.append(np.array(gen.last_residual_history))
            u = gen._workspace.download("sol", hi, 0)
            gen.close()
            if fast:
                u_fast = u
            else:
                assert np.array_equal(u, u_fast)
        assert hists[0].shape == hists[1].shape
        assert np.allclose(hists[0], hists[1], rtol=1e-12, atol=0)
        program = lowering.lower(expression, lo, hi)
        h, hist, rho, n = driver.solve(problem, program)
        assert len(hist) == len(hists[0])
        assert np.all(np.abs(hists[0] - np.array(hist)) <= 1e-10 * np.abs(hist))
        assert np.array_equal(u_fast, h.sol[hi][0])


ENGINE_VARIANTS = [
    {},                                                             # defaults: cluster tail (2D), lean, deferred zero fill
    {"MG_TAIL_CTAS": "1", "MG_TAIL_LEVEL": "6"},                    # single-CTA tail
    {"MG_TAIL_CTAS": "8", "MG_TAIL_LEVEL": "7", "MG_TAIL_SOLO": "3"},
    {"MG_TAIL_LEAN": "0"},                                          # full interpreter
    {"MG_LAZY_ZERO": "0"},                                          # zero fills issued where they stand
]


@pytest.mark.parametrize("dim", [2, 3])
def test_engine_configurations_are_bitwise_equivalent(dim, monkeypatch):
    """Tail kernel shape (single CTA / cluster / lean / full) and deferred zero fills are scheduling
    choices (read from the environment when a workspace is created): the converged solution and the
    residual history must not change by a single bit."""
    from evostencils_b200.generator import B200ProgramGenerator
    problem, lo, hi = (problems.POISSON_2D, 3, 10) if dim == 2 else (problems.POISSON_3D, 2, 6)
    grammar = Grammar(problem, lo, hi)
    expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 21 if dim == 2 else 23, "red_black"))
    reference = None
    for variant in ENGINE_VARIANTS:
        for name in ("MG_TAIL_CTAS", "MG_TAIL_LEVEL", "MG_TAIL_SOLO", "MG_TAIL_LEAN", "MG_LAZY_ZERO"):
            monkeypatch.delenv(name, raising=False)
        for name, value in variant.items():
            monkeypatch.setenv(name, value)
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
        gen.initialize_code_generation(lo, hi)
        gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
        got = (np.array(gen.last_residual_history), gen._workspace.download("sol", hi, 0))
        gen.close()
        if reference is None:
            reference = got
        else:
            assert np.array_equal(got[0], reference[0]), variant
            assert np.array_equal(got[1], reference[1]), variant


@pytest.mark.parametrize("block", [(1, 2), (1, 3), (1, 4), (2, 1), (2, 2), (3, 1), (4, 1)])
def test_fast_tile_smoother_bit_identical_to_generic(block):
    """Every block shape of the config-2 grammar (maximum_local_system_size = 4): the node-parallel
    tile kernel equals the generic one-thread-per-tile kernel bit for bit, and the oracle to 1e-12."""
    problem = problems.POISSON_2D
    lo, hi = 4, 9
    results = {}
    for fast in (True, False):
        ws, h = make_pair(problem, lo, hi, seed=11)
        ws.set_fast_kernels(fast)
        for level in (hi, hi - 3, lo):
            stmts = smoother_statements(problem, level, "block", ir.Single, omega=0.7, block=(block,))
            for sweep in range(2):
                if fast:
                    run_both(ws, h, stmts, problem)
                else:
                    ws.exec_ops(stmts)
            results[(fast, level)] = ws.download("sol", level, 0)
            if fast:
                compare(ws, h, "sol", level, False, f"fast block {block} level {level}")
        ws.close()
    for level in (hi, hi - 3, lo):
        assert np.array_equal(results[(True, level)], results[(False, level)]), (block, level)


@pytest.mark.parametrize("k", [5.0, 10.0, 80.0])
def test_helmholtz_preconditioned_bicgstab(k):
    """Outer BiCGStab with the cycle as preconditioner (Helmholtz/...exa3:144-200) on the GPU against
    the oracle: identical iteration count, per-iteration norms <= 1e-10 (`north_star`), the solution
    bit-identical.  BiCGStab amplifies rounding differences, so this only holds because oracle and
    engine share every operand order (complex product / quotient / root formulas, dot trees); k = 80
    is the reference's wave number (hundreds of iterations).  Levels 3..7 of the configuration:
    tests/test_gpu_config_sizes.py."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem, lo, hi = problems.HELMHOLTZ_2D, 3, 6
    grammar = Grammar(problem, lo, hi, parameters={"k": k})
    expression, _ = grammar.compile(v_cycle_string(hi - lo, 2, 1, 10, "red_black"))
    program = lowering.lower(expression, lo, hi)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.initialize_code_generation(lo, hi)
    t, rho, n = gen.evaluate_program(program, {"k": k}, evaluation_samples=2)
    h, hist, rho_ref, n_ref = driver.solve(problem, program, params={"k": k})
    got, want = np.array(gen.last_residual_history), np.array(hist)
    assert n == n_ref and n < 1e99
    assert got.shape == want.shape
    assert np.all(np.abs(got - want) <= 1e-10 * want), np.max(np.abs(got - want) / want)
    assert abs(rho - rho_ref) <= 1e-10 * rho_ref
    assert got[-1] < 1e-7 * got[0]
    gen.close()


@pytest.mark.parametrize("hi,one_pass", [(6, False), (7, False), (6, True), (7, True)])
def test_fast_3d_stages_bit_identical(hi, one_pass, monkeypatch):
    """2.5-D blocked 7-point kernels (levels >= 6 of 3D Poisson) against the oracle and the generic kernels;
    `one_pass`: RB-GS through k_rbgs3d (by default only from 257^3 up) instead of two colour passes."""
    from oracle import driver, mg_oracle
    monkeypatch.setenv("MG_RBGS3D_MIN_N", "65" if one_pass else "100000")
    monkeypatch.setenv("MG_RESTRICT3D_MIN_N", "65" if one_pass else "100000")     # k_resrestrict3d likewise
    problem, lo = problems.POISSON_3D, hi - 1
    ws, h = make_pair(problem, lo, hi, seed=13)
    run_both(ws, h, [lowering.Op("residual", hi, dst="res")], problem)
    compare(ws, h, "res", hi, True, "fast 3D residual")
    got, want = ws.residual_norm(hi), mg_oracle.residual_norm(h, hi)
    assert abs(got - want) <= 1e-12 * want
    for partitioning in (ir.Single, ir.RedBlack):
        stmts = smoother_statements(problem, hi, "collective", partitioning, omega=1.25)
        for sweep in range(3):
            run_both(ws, h, stmts, problem)
        compare(ws, h, "sol", hi, True, f"fast 3D smoother {partitioning.get_name()}")
    ops = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))
    ws.exec_ops(ops, fused=True)
    compare(ws, h, "rhs", hi - 1, True, "3D residual_restrict (fast residual + restriction)")
    run_both(ws, h, [lowering.Op("prolong_add", hi, omega=0.55, src="sol", dst="sol")], problem)
    compare(ws, h, "sol", hi, True, "fast 3D prolong_add")
    ws.close()


@pytest.mark.parametrize("omega", [1.15, 0.8])
def test_one_pass_3d_rbgs_bit_identical_to_generic_at_257(omega):
    """k_rbgs3d (TMA-staged planes, red half-sweep in shared memory) at 257^3 -- 5 x 33 x 9 tiles with
    clipped tiles on every side and random Dirichlet values -- against the generic colour passes."""
    from evostencils_b200.engine import Workspace
    problem, hi = problems.POISSON_3D, 8
    n = 2 ** hi + 1
    rng = np.random.default_rng(5)
    u0, f0 = rng.standard_normal((n, n, n)), rng.standard_normal((n, n, n))
    stmts = smoother_statements(problem, hi, "collective", ir.RedBlack, omega=omega)
    nc = 2 ** (hi - 1) + 1
    e0 = np.zeros((nc, nc, nc))
    e0[1:-1, 1:-1, 1:-1] = rng.standard_normal((nc - 2, nc - 2, nc - 2))
    out, coarse, corrected = [], [], []
    for fast in (True, False):
        ws = Workspace(3, hi - 1, hi, 1, "f64", 0)
        for level in (hi - 1, hi):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.set_fast_kernels(fast)
        ws.upload("sol", hi, 0, u0)
        ws.upload("rhs", hi, 0, f0)
        for sweep in range(2):
            ws.exec_ops(stmts)
        out.append(ws.download("sol", hi, 0))
        # fused residual + restriction (k_resrestrict3d: 5 x 32 x 4 tiles of coarse nodes) on the smoothed field
        ws.exec_ops([lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")], fused=True)
        coarse.append(ws.download("rhs", hi - 1, 0))
        # prolongation + correction with 8 rows per thread (the variant used from 257^3 up)
        ws.upload("sol", hi - 1, 0, e0)
        ws.exec_ops([lowering.Op("prolong_add", hi, omega=0.7, src="sol", dst="sol")])
        corrected.append(ws.download("sol", hi, 0))
        ws.close()
    bad = np.argwhere(out[0] != out[1])
    assert bad.size == 0, f"{len(bad)} mismatching nodes, first {bad[:4].tolist()}"
    assert not np.array_equal(out[0], u0)
    bad = np.argwhere(coarse[0][1:-1, 1:-1, 1:-1] != coarse[1][1:-1, 1:-1, 1:-1])
    assert bad.size == 0, f"restriction: {len(bad)} mismatching coarse nodes, first {bad[:4].tolist()}"
    assert np.abs(coarse[0][1:-1, 1:-1, 1:-1]).max() > 0
    bad = np.argwhere(corrected[0] != corrected[1])
    assert bad.size == 0, f"prolongation: {len(bad)} mismatching nodes, first {bad[:4].tolist()}"
    assert not np.array_equal(corrected[0], out[0])


def test_concurrent_evaluations_are_bitwise_those_of_sequential_ones():
    """Two generators (workspace + stream + host thread each) sharing a list of individuals
    (population.run_concurrently, the bench's `--streams 2`) give exactly the results and residual
    histories of one generator evaluating them one after the other."""
    from evostencils_b200 import population
    from evostencils_b200.generator import B200ProgramGenerator
    problem, lo, hi = problems.POISSON_2D, 5, 10
    grammar = Grammar(problem, lo, hi)
    strings = [v_cycle_string(hi - lo, 2, 1, 21, "red_black"), v_cycle_string(hi - lo, 2, 2, 14, "single"),
               v_cycle_string(hi - lo, 1, 1, 23, "red_black"), v_cycle_string(hi - lo, 3, 3, 12, "single"),
               v_cycle_string(hi - lo, 2, 0, 21, "red_black"), v_cycle_string(hi - lo, 1, 2, 16, "single")]

    def evaluator(gen, histories):
        def evaluate(string):
            expression, _ = grammar.compile(string)
            t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
            histories[string] = list(gen.last_residual_history)
            return (rho, n)
        return evaluate

    runs = []
    for n_gens in (1, 2):
        gens = [B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, engine_min_level=2)
                for _ in range(n_gens)]
        histories = {}
        got = population.run_concurrently([evaluator(g, histories) for g in gens], range(len(strings)), strings)
        runs.append(([got[j] for j in range(len(strings))], histories))
        for g in gens:
            g.close()
    assert runs[0][0] == runs[1][0]
    assert runs[0][1] == runs[1][1]
    assert all(len(h) > 3 for h in runs[0][1].values())

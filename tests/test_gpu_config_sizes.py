"""GPU parity at the sizes BASELINE.json's configurations name (through the C ABI).

  * config 2: level 12 (4097^2) -- every finest-level stage kernel at its bench geometry, including
    the `cp.async.bulk.prefetch.L2` paths that are only enabled for n > 2049 and every block shape of
    the config's grammar, against the numpy oracle; three cycles of population individuals (the two
    reference cycles and two block-Jacobi individuals, engine levels 2..12 like the bench) against the
    C/OpenMP restatement of the oracle (bit-identical to the numpy oracle: tests/test_oracle_c.py);
  * config 3: the TMA-staged 3D kernels at their default thresholds against the numpy oracle at 257^3
    and against the C restatement at 513^3;
  * config 4: Helmholtz on levels 3..7 at k = 80, identical iteration count and per-iteration norms
    <= 1e-10 (the oracle states the complex arithmetic and the reduction trees of the engine);
  * config 5: elasticity on levels 4..8, RB-GS baseline and the ((2,1),(2,1)) block-Jacobi individual;
  * f4: level generalisation + PDE-parameter sweep through `reinitialize` and `generate_and_evaluate`.
"""
import os

import numpy as np
import pytest

from evostencils_b200 import ir, lowering, problems
from evostencils_b200.grammar import Grammar, v_cycle_string

from test_gpu_parity import compare, make_pair, run_both, smoother_statements

pytestmark = pytest.mark.gpu

POPULATION = os.path.join(os.path.dirname(__file__), "..", "evostencils_b200", "data", "population_2d_poisson_l7_12.txt")


def exact_equal(got, want, what):
    bad = np.argwhere(got != want)
    assert bad.size == 0, f"{what}: {len(bad)} mismatching nodes, first {bad[:3].tolist()}, " \
                          f"max abs diff {np.max(np.abs(got - want))}"


# ---- config 2: 4097^2 ---------------------------------------------------------------------------------

def test_level12_stage_kernels_against_the_numpy_oracle():
    """One launch of every finest-level stage kernel at 4097^2 (grids of 16 x 1024 CTAs, R = 4 rows,
    L2 prefetch 20 bands ahead) -- bit-identical to the numpy oracle."""
    from oracle import driver, mg_oracle
    problem, lo, hi = problems.POISSON_2D, 11, 12
    ws, h = make_pair(problem, lo, hi, seed=21)
    run_both(ws, h, [lowering.Op("residual", hi, dst="res")], problem)
    compare(ws, h, "res", hi, True, "residual 4097^2")
    got, want = ws.residual_norm(hi), mg_oracle.residual_norm(h, hi)
    assert abs(got - want) <= 1e-12 * want
    for partitioning in (ir.Single, ir.RedBlack):
        stmts = smoother_statements(problem, hi, "collective", partitioning, omega=1.15)
        run_both(ws, h, stmts, problem)
        compare(ws, h, "sol", hi, True, f"{partitioning.get_name()} sweep 4097^2")
    # the zero-input variant (deferred fill consumed by the smoother) on both levels
    for level in (hi, lo):
        ops = [lowering.Op("zero", level, dst="sol")] + smoother_statements(problem, level, "collective", ir.RedBlack, 1.15)
        run_both(ws, h, ops, problem)
        compare(ws, h, "sol", level, True, f"zero-input RB-GS level {level}")
        h.sol[level][0][...] = np.random.default_rng(level).standard_normal(h.sol[level][0].shape)
        if level < hi:
            h.sol[level][0][0, :] = h.sol[level][0][-1, :] = 0
            h.sol[level][0][:, 0] = h.sol[level][0][:, -1] = 0
        ws.upload("sol", level, 0, h.sol[level][0])
    ops = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))
    ws.exec_ops(ops, fused=True)
    compare(ws, h, "rhs", hi - 1, True, "fused residual + restriction 4097^2 -> 2049^2")
    run_both(ws, h, [lowering.Op("prolong_add", hi, omega=0.55, src="sol", dst="sol")], problem)
    compare(ws, h, "sol", hi, True, "prolongation + correction 2049^2 -> 4097^2")
    ws.close()


@pytest.mark.parametrize("block", [(1, 2), (1, 3), (1, 4), (2, 1), (2, 2), (3, 1), (4, 1)])
def test_level12_block_jacobi_against_the_numpy_oracle(block):
    problem, lo, hi = problems.POISSON_2D, 11, 12
    ws, h = make_pair(problem, lo, hi, seed=22)
    stmts = smoother_statements(problem, hi, "block", ir.Single, omega=0.7, block=(block,))
    run_both(ws, h, stmts, problem)
    compare(ws, h, "sol", hi, False, f"block {block} 4097^2")
    ws.close()


def bench_inputs(level):
    n = 2 ** level + 1
    rng = np.random.default_rng(1234)
    f = np.zeros((n, n))
    f[1:-1, 1:-1] = rng.standard_normal((n - 2, n - 2))
    return [np.zeros((n, n))], [f]


def population_strings():
    with open(POPULATION) as fh:
        strings = [line.strip() for line in fh if line.strip() and not line.startswith("#")]
    finest_block = [s for s in strings if "collective_block_jacobi_0(" in s]
    return [strings[0], strings[1]] + finest_block[:2]


def test_level12_population_cycles_against_the_c_port():
    """Three cycles of four individuals of the stored population exactly as the bench runs them (grammar
    levels 7..12, default cycle below, engine levels 2..12, N(0,1) right-hand side)."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import cport
    problem, lo, hi, elo = problems.POISSON_2D, 7, 12, 2
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    us, fs = bench_inputs(hi)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, engine_min_level=elo)
    gen.set_inputs(rhs=fs, u0=us)
    cws = cport.workspace(2, elo, hi)
    for level in range(elo, hi + 1):
        cws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    cws.set_coarse_solver("CG", problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    strings = population_strings()
    assert len(strings) == 4
    for k, string in enumerate(strings):
        program = lowering.lower(grammar.compile(string)[0], lo, hi)
        gen.evaluate_program(program, {}, evaluation_samples=1, max_iterations=3)
        got = np.array(gen.last_residual_history)
        cws.upload("sol", hi, 0, us[0])
        cws.upload("rhs", hi, 0, fs[0])
        cws.clear_programs()
        cycle = cws.build_cycle(program.ops, lambda level: lowering.default_cycle_ops(problem, level, elo), fused=False)
        want, _, _ = cws.solve(cycle, problem.solver.target_reduction, 3)
        want = np.array(want)
        assert got.shape == want.shape == (4,), (k, got, want)
        assert np.all(np.abs(got - want) <= 1e-10 * np.abs(want)), (k, got, want)
        u, u_ref = gen._workspace.download("sol", hi, 0), cws.download("sol", hi, 0)
        if "collective_block_jacobi" in string:
            assert np.max(np.abs(u - u_ref)) <= 1e-12 * np.max(np.abs(u_ref)), k
        else:
            exact_equal(u, u_ref, f"individual {k} after 3 cycles")
    cws.close()
    gen.close()


# ---- config 3: 257^3 (numpy oracle) and 513^3 (C restatement) ---------------------------------------------

def test_level8_3d_stage_kernels_against_the_numpy_oracle():
    """k_rbgs3d / k_resrestrict3d / k_prolong3d<..., 8> at their default thresholds (257^3)."""
    from oracle import driver, mg_oracle
    problem, lo, hi = problems.POISSON_3D, 7, 8
    ws, h = make_pair(problem, lo, hi, seed=23)
    stmts = smoother_statements(problem, hi, "collective", ir.RedBlack, omega=1.25)
    run_both(ws, h, stmts, problem)
    compare(ws, h, "sol", hi, True, "one-pass RB-GS 257^3")
    ops = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    mg_oracle.run_ops(h, ops, driver.settings_for(problem))
    ws.exec_ops(ops, fused=True)
    compare(ws, h, "rhs", hi - 1, True, "fused residual + restriction 257^3 -> 129^3")
    run_both(ws, h, [lowering.Op("prolong_add", hi, omega=0.7, src="sol", dst="sol")], problem)
    compare(ws, h, "sol", hi, True, "prolongation + correction 129^3 -> 257^3")
    ws.close()


def test_level9_3d_stage_kernels_against_the_c_port():
    """The same three kernels at config 3's size, 513^3 (1.08 GB per field), against the C/OpenMP
    restatement of the oracle; bit-identical."""
    from evostencils_b200.engine import Workspace
    from oracle import cport
    problem, lo, hi = problems.POISSON_3D, 8, 9
    n, nc = 2 ** hi + 1, 2 ** lo + 1
    rng = np.random.default_rng(24)
    u0 = rng.standard_normal((n, n, n))
    f0 = rng.standard_normal((n, n, n))
    e0 = np.zeros((nc, nc, nc))
    e0[1:-1, 1:-1, 1:-1] = rng.standard_normal((nc - 2,) * 3)
    stmts = smoother_statements(problem, hi, "collective", ir.RedBlack, omega=1.25)
    fused = [lowering.Op("residual", hi, dst="res"), lowering.Op("restrict", hi - 1, src="res", dst="rhs")]
    prolong = [lowering.Op("prolong_add", hi, omega=0.7, src="sol", dst="sol")]
    results = []
    for make in (lambda: Workspace(3, lo, hi, 1, "f64", 0), lambda: cport.workspace(3, lo, hi)):
        ws = make()
        for level in (lo, hi):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.upload("sol", hi, 0, u0)
        ws.upload("rhs", hi, 0, f0)
        ws.upload("sol", lo, 0, e0)
        out = []
        ws.exec_ops(stmts)
        out.append(ws.download("sol", hi, 0))
        ws.exec_ops(fused, fused=ws.lib is not cport.load())
        out.append(ws.download("rhs", lo, 0)[1:-1, 1:-1, 1:-1].copy())
        ws.exec_ops(prolong)
        out.append(ws.download("sol", hi, 0))
        ws.close()
        results.append(out)
    for name, got, want in zip(("one-pass RB-GS 513^3", "residual + restriction 513^3", "prolongation 513^3"), *results):
        exact_equal(got, want, name)
    assert not np.array_equal(results[0][0], u0) and np.abs(results[0][1]).max() > 0


# ---- config 4: Helmholtz 3..7, k = 80 -------------------------------------------------------------------------

def helmholtz_pair(lo, hi, k, string=None, max_iterations=None, gen=None):
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem = problems.HELMHOLTZ_2D
    grammar = Grammar(problem, lo, hi, parameters={"k": k})
    expression, _ = grammar.compile(string or v_cycle_string(hi - lo, 2, 1, 10, "red_black"))
    program = lowering.lower(expression, lo, hi)
    own = gen is None
    if own:
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    t, rho, n = gen.evaluate_program(program, {"k": k}, evaluation_samples=1, max_iterations=max_iterations)
    got = np.array(gen.last_residual_history)
    if own:
        gen.close()
    _, hist, rho_ref, n_ref = driver.solve(problem, program, params={"k": k}, max_iterations=max_iterations)
    return got, np.array(hist), (rho, n), (rho_ref, n_ref)


def assert_histories_identical(got, want, tol=1e-10):
    assert got.shape == want.shape, (len(got), len(want))
    rel = np.abs(got - want) / np.abs(want)
    assert np.all(rel <= tol), (int(np.argmax(rel > tol)), float(rel.max()))


def test_helmholtz_config4_identical_iteration_count():
    """`north_star`: identical iteration count and per-iteration norms <= 1e-10 -- at the reference's
    configuration (levels 3..7, k = 80, V(2,1) RB-GS w = 0.6 preconditioner: ~600 BiCGStab iterations)."""
    got, want, (rho, n), (rho_ref, n_ref) = helmholtz_pair(3, 7, 80.0)
    assert n == n_ref and n < 1e99
    assert_histories_identical(got, want)
    assert abs(rho - rho_ref) <= 1e-10 * rho_ref


@pytest.mark.parametrize("k", [5.0, 20.0])
def test_helmholtz_small_wave_numbers(k):
    got, want, (rho, n), (rho_ref, n_ref) = helmholtz_pair(3, 6, k)
    assert n == n_ref
    assert_histories_identical(got, want)


def test_helmholtz_jacobi_preconditioner_after_another_individual():
    """An odd number of Jacobi (`single`) sweeps on the finest level leaves the ping-pong state
    flipped when the iteration graph is captured; the initial guess must still be the uploaded one,
    whatever individual ran before on the workspace."""
    from evostencils_b200.generator import B200ProgramGenerator
    problem, lo, hi, k = problems.HELMHOLTZ_2D, 3, 5, 5.0
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    first = helmholtz_pair(lo, hi, k, gen=gen, max_iterations=20)
    assert_histories_identical(first[0], first[1])
    jacobi = v_cycle_string(hi - lo, 2, 1, 10, "single")            # 3 sweeps on the finest level
    got, want, _, _ = helmholtz_pair(lo, hi, k, string=jacobi, gen=gen, max_iterations=20)
    assert_histories_identical(got, want)
    gen.close()


# ---- config 5: elasticity 4..8 ---------------------------------------------------------------------------------

@pytest.mark.parametrize("kind", ["rbgs", "block"])
def test_elasticity_config5(kind):
    """Levels 4..8 (LinearElasticity/...knowledge:3-4): the `.exa3` baseline V(2,1) collective RB-GS
    w = 1.25 and the collective block-Jacobi individual with block shape ((2,1),(2,1))."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem, lo, hi = problems.LINEAR_ELASTICITY_2D, 4, 8
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    if kind == "rbgs":
        string, max_its = v_cycle_string(hi - lo, 2, 1, 23, "red_black"), None
    else:
        string = v_cycle_string(hi - lo, 2, 1, 10, "((2, 1), (2, 1))", smoother="collective_block_jacobi")
        max_its = 25
    program = lowering.lower(grammar.compile(string)[0], lo, hi)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.evaluate_program(program, {}, evaluation_samples=1, max_iterations=max_its)
    got = np.array(gen.last_residual_history)
    h, hist, rho_ref, n_ref = driver.solve(problem, program, max_iterations=max_its)
    assert_histories_identical(got, np.array(hist))
    for f in range(2):
        u = gen._workspace.download("sol", hi, f)
        if kind == "rbgs":
            exact_equal(u, h.sol[hi][f], f"field {f}")
        else:
            assert np.max(np.abs(u - h.sol[hi][f])) <= 1e-12 * np.max(np.abs(h.sol[hi][f]))
    gen.close()


# ---- f4: level generalisation + parameter sweep ----------------------------------------------------------------

def test_level_generalisation_and_wave_number_sweep():
    """optimization/program.py:110-146, 515-540: levels + 1, k doubled, the same individuals
    re-evaluated; inside generate_and_evaluate the wave number runs k, 2k, 4k on the same grid
    (exastencils.py:518-532).  Every one of the solves against the oracle."""
    from evostencils_b200.generator import B200ProgramGenerator
    from oracle import driver
    problem = problems.HELMHOLTZ_2D
    lo, hi, k = 3, 5, 5.0
    string = v_cycle_string(hi - lo, 2, 1, 10, "red_black")
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)

    def evaluate(lo, hi, k):
        grammar = Grammar(problem, lo, hi, parameters={"k": k})
        expression, _ = grammar.compile(string)
        gen.initialize_code_generation(lo, hi)
        storages = gen.generate_storage(lo, hi, gen.finest_grid)
        result = gen.generate_and_evaluate(expression, storages, lo, hi, "", evaluation_samples=1,
                                           global_variable_values={"k": k})
        assert len(gen.last_sweep) == 3 and [s[0]["k"] for s in gen.last_sweep] == [k, 2 * k, 4 * k]
        total = [0.0, 0.0, 0.0]
        for params, t, rho, n, history in gen.last_sweep:
            override = lambda level, p=params: problem.system_operator(level, 0, p)
            program = lowering.lower(expression, lo, hi, override)
            _, want, rho_ref, n_ref = driver.solve(problem, program, params=params)
            assert n == n_ref, params
            assert_histories_identical(np.array(history), np.array(want))
            total = [total[0] + t, total[1] + rho, total[2] + n]
        assert result == pytest.approx(tuple(v / 3 for v in total))

    evaluate(lo, hi, k)
    gen.reinitialize(lo + 1, hi + 1, {"k": 2 * k})
    assert (gen.min_level, gen.max_level) == (lo + 1, hi + 1)
    assert gen.finest_grid[0].level == hi + 1 and gen.finest_grid[0].size == (2 ** (hi + 1),) * 2
    evaluate(lo + 1, hi + 1, 2 * k)
    gen.close()


def test_evaluation_timeout_and_unexecutable_individuals_give_the_sentinel():
    """exastencils.py:424-431: a run that exceeds `timeout_evaluate`, like any per-individual failure,
    scores (infinity,)*3 instead of raising."""
    from evostencils_b200.generator import B200ProgramGenerator
    problem, lo, hi = problems.POISSON_2D, 5, 11
    grammar = Grammar(problem, lo, hi)
    slow = v_cycle_string(hi - lo, 1, 0, 2, "single")            # w = 0.2 Jacobi: far from converging in 100 cycles
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, evaluation_timeout=1e-4)
    assert gen.generate_and_evaluate(grammar.compile(slow)[0], None, lo, hi, "", evaluation_samples=1) == (1e100,) * 3
    assert "time limit" in gen.last_error
    gen.timeout_evaluate = 300
    t, rho, n = gen.generate_and_evaluate(grammar.compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))[0], None, lo, hi,
                                          "", evaluation_samples=1)
    assert n < 20 and rho < 0.2
    # a Richardson-type coarse correction (residual without a smoother: ill-typed, but `eval` accepts it)
    # lowers to an op the engine does not execute -> sentinel, no exception
    bad = ("update_with_coarse_grid_correction_0(18, P_1, "
           "(coarsening_0(A_1, zero_1, R_0, residual_0(u_and_f)), None))")
    assert gen.generate_and_evaluate(grammar.compile(bad)[0], None, lo, hi, "", evaluation_samples=1) == (1e100,) * 3
    assert "richardson" in gen.last_error
    gen.close()

"""Nonlinear FAS path (example_problems/FAS_2D_Basic): oracle self-checks on the
CPU, parity of the CUDA stages and of whole solves on the GPU.  exp() differs between libm and the
device by an ulp, so fields are compared to 1e-12 relative, residual histories to 1e-10 (north_star)."""
import numpy as np
import pytest

from evostencils_b200 import fas


def test_cycle_statement_order_is_that_of_the_generated_listing():
    """FAS_2D_Basic.exa4:210-269: V(2,2) over levels 5, 4, 3 with CGS = 30 sweeps."""
    got = [(op.kind, op.level, op.sweeps) for op in fas.v_cycle_ops(5, 3)]
    assert got == [("smooth", 5, 2), ("residual", 5, 1), ("restrict", 4, 1),
                   ("smooth", 4, 2), ("residual", 4, 1), ("restrict", 3, 1),
                   ("smooth", 3, 30),
                   ("subtract", 3, 1), ("prolong_add", 4, 1), ("smooth", 4, 2),
                   ("subtract", 4, 1), ("prolong_add", 5, 1), ("smooth", 5, 2)]


def parse_layer4(text):
    """Statement sequence of a cycle function emitted by the reference's ProgramGeneratorFAS:
    smooth -> ("smooth", level, sweeps, omega, newton_steps, colored) with omega None for calls of the
    template's Smoother function; the others (kind, level, factor or 1)."""
    import re
    out = []
    updates, in_colour = None, False

    def add_smooth(level, omega, newton_steps, colored):
        last = out[-1] if out else None
        if last and last[0] == "smooth" and (last[1], last[3], last[4], last[5]) == (level, omega, newton_steps, colored):
            out[-1] = ("smooth", level, last[2] + 1, omega, newton_steps, colored)
        else:
            out.append(("smooth", level, 1, omega, newton_steps, colored))

    for line in text.splitlines():
        line = line.strip()
        m = re.match(r"Smoother@(\d+) \(", line)
        if m:
            add_smooth(int(m.group(1)), None, 1, False)
            continue
        m = re.match(r"CGS@(\d+) \(", line)
        if m:
            out.append(("cgs", int(m.group(1)), 1))
            continue
        if line.startswith("color with"):
            in_colour = True
            continue
        m = re.match(r"loop over Solution@(\d+) \{", line)
        if m:
            updates = []
            continue
        m = re.match(r"Solution@(\d+) \+= ([0-9.e-]+) \* \( \( RHS@\1 - \( \( Laplace@\1 \* Solution@\1 \) \+ "
                     r"\( gamSten@\1 \* Solution@\1 \) \) \) / (.*) \)$", line)
        if m:
            den = m.group(3)
            assert den in (f"Laplace@{m.group(1)}:[0,0]",
                           f"( Laplace@{m.group(1)}:[0,0] + gam*(Solution@{m.group(1)} + 1)*exp(Solution@{m.group(1)}) )"), den
            updates.append((int(m.group(1)), float(m.group(2)), "gam" in den))
            continue
        if line == "}" and updates:
            level, omega, newton = updates[0]
            assert all(u == updates[0] for u in updates)
            add_smooth(level, omega, len(updates) if newton else 0, in_colour)
            updates = None
            continue
        if line == "}" and in_colour and updates is None:
            in_colour = False                                       # end of the colour block (after `apply bc`)
            continue
        if line == "}":
            updates = None
            continue
        m = re.match(r"Residual@(\d+) = \( RHS@\1 - \( \( Laplace@\1 \* Solution@\1 \) \+ \( gamSten@\1 \* Solution@\1 \) \) \)", line)
        if m:
            out.append(("residual", int(m.group(1)), 1))
            continue
        m = re.match(r"Approximation@(\d+) = \( RestrictionNode@(\d+) \* Solution@\d+ \)", line)
        if m:
            assert int(m.group(2)) == int(m.group(1)) + 1
            out.append(("restrict_solution", int(m.group(1)), 1))
            continue
        m = re.match(r"RHS@(\d+) = \( \( RestrictionNode@(\d+) \* Residual@\2 \) \+ \( \( Laplace@\1 \* Approximation@\1 \) "
                     r"\+ \( gamSten@\1 \* Approximation@\1 \) \) \)", line)
        if m:
            assert out[-1] == ("restrict_solution", int(m.group(1)), 1)      # the two loops of one `restrict` op
            out[-1] = ("restrict", int(m.group(1)), 1)
            continue
        m = re.match(r"Solution@(\d+) -= Approximation@\1", line)
        if m:
            out.append(("subtract", int(m.group(1)), 1))
            continue
        m = re.match(r"Solution@(\d+) \+= \( ([0-9.e-]+) \* \( CorrectionNode@(\d+) \* Solution@\3 \) \)", line)
        if m:
            assert int(m.group(3)) == int(m.group(1)) - 1
            out.append(("prolong_add", int(m.group(1)), float(m.group(2))))
            continue
        assert line == "" or line.startswith(("Function", "loop over", "apply bc", "communicate", "advance", "Variable",
                                              "Solution<next>@", "Solution@", "( i0 + i1 ) % 2")), line
    return out


def as_statements(ops, template_smoother):
    got = []
    for op in ops:
        if op.kind == "smooth" and op.cgs:
            got.append(("cgs", op.level, 1))                        # CGS@coarsest = 30 sweeps of the smoother
        elif op.kind == "smooth":
            got.append(("smooth", op.level, op.sweeps, None if template_smoother else op.omega,
                        1 if template_smoother else op.newton_steps, False if template_smoother else op.colored))
        elif op.kind == "prolong_add":
            got.append(("prolong_add", op.level, op.omega))
        else:
            got.append((op.kind, op.level, 1))
    return got


def load_golden():
    import json
    import os
    return json.load(open(os.path.join(os.path.dirname(__file__), "golden", "fas_golden.json")))


def test_cycle_statement_order_is_that_of_the_reference_emitter():
    """tests/golden/fas_golden.json: Layer-4 text emitted by the unmodified reference
    (`ProgramGeneratorFAS.generate_cycle_function`) for FAS V-cycles written in its FAS grammar."""
    golden = load_golden()
    cases = {k: v for k, v in golden.items() if "pre" in v}
    assert len(cases) >= 4
    for name, case in cases.items():
        want = parse_layer4(case["layer4"])
        factors = {s[2] for s in want if s[0] == "prolong_add"}
        assert len(factors) == 1, name
        ops = fas.v_cycle_ops(case["max_level"], case["min_level"], case["pre"], case["post"], cgc_omega=factors.pop())
        assert as_statements(ops, True) == want, name


def test_lowering_of_fas_grammar_strings_matches_the_reference_emitter():
    """Grammar string -> (our FAS productions) -> IR graph -> `fas.lower` against the statements of the
    listing the reference emits for the same string: template smoother calls and inline Jacobi-Picard /
    Jacobi-Newton / red-black smoothers, V and W-like shapes."""
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    golden = load_golden()
    assert sum(1 for c in golden.values() if not c["template_smoother"]) >= 3
    for name, case in golden.items():
        grammar = Grammar(problems.FAS_2D_BASIC, case["min_level"], case["max_level"], FAS=True)
        expression, _ = grammar.compile(case["grammar"])
        ops = fas.lower(expression, template_smoother=case["template_smoother"])
        assert as_statements(ops, case["template_smoother"]) == parse_layer4(case["layer4"]), name
        assert fas.lower(grammar.compile(case["grammar"])[0], case["template_smoother"]) == ops      # graph not consumed


def test_layer4_export_is_byte_identical_to_the_reference_emitter():
    """`generate_cycle_function` of the generator mirror returns the text the reference returns."""
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    for name, case in load_golden().items():
        gen = fas.B200ProgramGeneratorFAS("FAS_2D_Basic", "Solution", "RHS", "Residual", "Approximation", "RestrictionNode",
                                          "CorrectionNode", "Laplace", "gamSten", "gen_mgCycle", "CGS",
                                          "Smoother" if case["template_smoother"] else None)
        expression, _ = Grammar(problems.FAS_2D_BASIC, case["min_level"], case["max_level"], FAS=True).compile(case["grammar"])
        assert gen.generate_cycle_function(expression) == case["layer4"], name


def test_lowering_accepts_the_reference_ir():
    """The drop-in boundary: graphs built by the reference's own grammar layer lower to the same
    statements (container only: needs /root/reference)."""
    import os
    import sys
    if not os.path.isdir("/root/reference/evostencils"):
        pytest.skip("reference checkout not present")
    make = os.path.join(os.path.dirname(__file__), "golden", "_make")
    sys.path[:0] = [make, "/root/reference"]
    cwd = os.getcwd()
    try:
        os.chdir("/root/reference")
        import importlib.util
        spec = importlib.util.spec_from_file_location("make_fas_golden", os.path.join(make, "make_fas_golden.py"))
        gen = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(gen)
        pg = gen.ProgramGeneratorFAS("FAS_2D_Basic", "Solution", "RHS", "Residual", "Approximation", "RestrictionNode",
                                     "CorrectionNode", "Laplace", "gamSten", "gen_mgCycle", "CGS", None)
        for name, case in load_golden().items():
            pset = gen.front_end(pg, case["max_level"] - case["min_level"])
            expression, _ = eval(case["grammar"], pset.context, {})
            ops = fas.lower(expression, template_smoother=case["template_smoother"])
            assert as_statements(ops, case["template_smoother"]) == parse_layer4(case["layer4"]), name
    finally:
        os.chdir(cwd)
        sys.path.remove(make)
        sys.path.remove("/root/reference")


def test_generator_mirror_has_the_reference_interface():
    """Attributes the optimizer reads from a ProgramGeneratorFAS (optimization/program.py:806-842;
    scripts that construct it) exist without touching the GPU."""
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    gen = fas.B200ProgramGeneratorFAS("FAS_2D_Basic", "Solution", "RHS", "Residual", "Approximation", "RestrictionNode",
                                      "CorrectionNode", "Laplace", "gamSten", "gen_mgCycle", "CGS", "Smoother")
    assert gen.uses_FAS and (gen.dimension, gen.min_level, gen.max_level) == (2, 6, 10)       # .knowledge file
    assert gen.generate_storage(6, 10, gen.finest_grid) == [] and gen.initialize_code_generation(6, 10) is None
    assert gen.finest_grid[0].level == 10 and gen.coarsening_factor == [(2, 2)]
    assert len(gen.equations) == 5 and len(gen.operators) == 15 and len(gen.fields) == 1
    assert {o.name for o in gen.operators} == {"Laplace", "RestrictionNode", "ProlongationNode"}   # exastencils_FAS.py:77-85
    expression, _ = Grammar(problems.FAS_2D_BASIC, 8, 10, FAS=True).compile(load_golden()["fas_v22_two_coarsenings"]["grammar"])
    assert gen.generate_cycle_function(expression) == load_golden()["fas_v22_two_coarsenings"]["layer4"]
    assert gen.lower(expression) == fas.v_cycle_ops(10, 8, 2, 2, cgc_omega=gen.lower(expression)[-2].omega)
    import torch
    if not torch.cuda.is_available():
        from evostencils_b200.engine import EngineError
        with pytest.raises(EngineError, match="no CPU fallback"):
            gen.generate_and_evaluate(expression, evaluation_samples=1)


def test_oracle_converges_to_the_discretisation_error():
    from oracle import fas_oracle as fo
    errors = []
    for hi in (5, 6, 7):
        u, history = fo.solve(hi, 3)
        assert history[-1] < 1e-10 * history[0] and len(history) - 1 < 20
        factors = [history[i + 1] / history[i] for i in range(len(history) - 1)]
        assert max(factors) < 0.25                                  # a working V(2,2) cycle
        errors.append(np.abs(u - fo.exact_solution(hi)).max())
    assert 3.7 < errors[0] / errors[1] < 4.3 and 3.7 < errors[1] / errors[2] < 4.3     # O(h^2)


def test_oracle_stages_are_consistent():
    from oracle import fas_oracle as fo
    rng = np.random.default_rng(3)
    level = 5
    n = fo.nodes(level)
    u = np.zeros((n, n))
    u[1:-1, 1:-1] = 0.1 * rng.standard_normal((n - 2, n - 2))
    f = fo.apply_operator(u, level, fo.GAMMA)
    assert fo.norm(fo.residual(u, f, level, fo.GAMMA)) == 0.0
    assert np.array_equal(fo.smooth(u, f, level, fo.GAMMA), u)      # a solution is a fixed point of the smoother
    # FAS coarse problem of an exact fine solution: the coarse solve returns the restricted solution,
    # the correction is zero
    approx = fo.restrict(u)
    f_c = fo.restrict(fo.residual(u, f, level, fo.GAMMA))
    f_c[1:-1, 1:-1] += fo.apply_operator(approx, level - 1, fo.GAMMA)[1:-1, 1:-1]
    assert np.array_equal(fo.smooth(approx, f_c, level - 1, fo.GAMMA), approx)
    # host-side problem definition = the oracle's
    assert np.array_equal(fas.right_hand_side(level), fo.right_hand_side(level))
    assert np.array_equal(fas.exact_solution(level), fo.exact_solution(level))


def test_c_port_of_the_fas_stages_matches_the_numpy_oracle():
    """oracle/mg_oracle_c.c carries the FAS statements too (the timed CPU figure of
    scripts/evaluate_fas_solver.py); same arithmetic as oracle/fas_oracle.py up to libm's exp."""
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    from oracle import cport
    from oracle import fas_oracle as fo
    golden = load_golden()
    for case, lo, hi in (("fas_inline_w_like", 3, 6), ("fas_inline_v_two_coarsenings", 4, 6), ("fas_v22_two_coarsenings", 3, 5)):
        expression, _ = Grammar(problems.FAS_2D_BASIC, lo, hi, FAS=True).compile(golden[case]["grammar"])
        ops = fas.lower(expression, template_smoother=golden[case]["template_smoother"])
        solver = cport.FasSolver(lo, hi)
        history, ms = solver.solve(ops)
        u_ref, want = fo.solve_ops(ops, hi, lo)
        assert len(history) == len(want), case
        assert np.all(np.abs(np.array(history) - np.array(want)) <= 1e-10 * np.array(want) + 1e-13 * want[0]), case
        assert np.max(np.abs(solver.solution() - u_ref)) <= 1e-10 * np.max(np.abs(u_ref)), case
        solver.close()


def close(got, want, tol=1e-12):
    scale = np.max(np.abs(want)) + 1e-300
    return np.max(np.abs(got - want)) <= tol * scale


@pytest.mark.gpu
@pytest.mark.parametrize("level", [5, 8])
def test_fas_stages_match_the_oracle(level):
    from oracle import fas_oracle as fo
    rng = np.random.default_rng(11)
    n = fo.nodes(level)
    u = np.zeros((n, n))
    u[1:-1, 1:-1] = 0.3 * rng.standard_normal((n - 2, n - 2))
    f = rng.standard_normal((n, n)) * float(4 ** level) * 0.1
    solver = fas.FasSolver(level - 1, level)
    ws = solver.ws
    ws.upload("sol", level, 0, u)
    ws.upload("rhs", level, 0, f)
    got_norm, want_norm = ws.fas_residual_norm(level, fo.GAMMA), fo.norm(fo.residual(u, f, level, fo.GAMMA))
    assert abs(got_norm - want_norm) <= 1e-12 * want_norm
    ws.fas_smooth(level, 0.8, fo.GAMMA, 3)
    want = u
    for _ in range(3):
        want = fo.smooth(want, f, level, fo.GAMMA, 0.8)
    assert close(ws.download("sol", level, 0), want), "Jacobi-Newton sweeps"
    ws.fas_residual(level, fo.GAMMA)
    r = fo.residual(want, f, level, fo.GAMMA)
    assert close(ws.download("res", level, 0), r), "residual"
    ws.fas_restrict(level - 1, fo.GAMMA)
    approx = fo.restrict(want)
    f_c = fo.restrict(r)
    f_c[1:-1, 1:-1] = f_c[1:-1, 1:-1] + fo.apply_operator(approx, level - 1, fo.GAMMA)[1:-1, 1:-1]
    assert close(ws.download("sol", level - 1, 0), approx), "restricted solution"
    assert close(ws.download("rhs", level - 1, 0)[1:-1, 1:-1], f_c[1:-1, 1:-1]), "FAS coarse right-hand side"
    ws.fas_smooth(level - 1, 0.8, fo.GAMMA, 2)
    u_c = fo.smooth(fo.smooth(approx, f_c, level - 1, fo.GAMMA, 0.8), f_c, level - 1, fo.GAMMA, 0.8)
    ws.fas_subtract_approximation(level - 1)
    assert close(ws.download("sol", level - 1, 0), u_c - approx, 1e-9), "correction"     # a difference of close values
    solver.run_cycle([fas.FasOp("prolong_add", level)])
    assert close(ws.download("sol", level, 0), fo.prolong_add(want, u_c - approx), 1e-11), "prolongation + correction"
    solver.close()


@pytest.mark.gpu
@pytest.mark.parametrize("lo,hi,pre,post,cgc", [(3, 5, 2, 2, 1.0), (3, 7, 2, 2, 1.0), (4, 8, 1, 1, 1.0), (3, 6, 1, 2, 0.7)])
def test_fas_solve_matches_the_oracle(lo, hi, pre, post, cgc):
    from oracle import fas_oracle as fo
    solver = fas.FasSolver(lo, hi)
    history, ms = solver.solve(fas.v_cycle_ops(hi, lo, pre, post, cgc_omega=cgc))
    u_ref, want = fo.solve(hi, lo, pre, post, cgc_omega=cgc)
    assert len(history) == len(want)                                # identical iteration count
    assert np.all(np.abs(np.array(history) - np.array(want)) <= 1e-10 * np.array(want) + 1e-13 * want[0])
    assert close(solver.solution(), u_ref, 1e-11)
    assert np.abs(solver.solution() - fas.exact_solution(hi)).max() < 2.0 * 4.0 ** (5 - hi) * 1e-3
    solver.close()


@pytest.mark.gpu
@pytest.mark.parametrize("newton_steps,colored,omega", [(2, False, 0.8), (0, False, 0.7), (1, True, 1.1), (0, True, 1.1), (3, True, 0.9)])
def test_fas_smoother_variants_match_the_oracle(newton_steps, colored, omega):
    """Jacobi-Picard, several local Newton steps and the red-black passes (exastencils_FAS.py:206-246)."""
    from oracle import fas_oracle as fo
    level = 7
    rng = np.random.default_rng(17)
    n = fo.nodes(level)
    u = np.zeros((n, n))
    u[1:-1, 1:-1] = 0.3 * rng.standard_normal((n - 2, n - 2))
    f = rng.standard_normal((n, n)) * float(4 ** level) * 0.1
    solver = fas.FasSolver(level - 1, level)
    solver.ws.upload("sol", level, 0, u)
    solver.ws.upload("rhs", level, 0, f)
    solver.ws.fas_smooth(level, omega, fo.GAMMA, 3, newton_steps, colored)
    want = u
    for _ in range(3):
        want = fo.smooth(want, f, level, fo.GAMMA, omega, newton_steps, colored)
    assert close(solver.ws.download("sol", level, 0), want)
    solver.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["fas_inline_picard_newton", "fas_inline_v_two_coarsenings", "fas_inline_w_like",
                                  "fas_v12_damped_correction"])
def test_lowered_fas_individuals_solve_like_the_oracle(case):
    """Grammar string -> IR -> statement list -> GPU solve, against the numpy execution of the same list."""
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    from oracle import fas_oracle as fo
    golden = load_golden()[case]
    depth = golden["max_level"] - golden["min_level"]
    hi, lo = 7, 7 - depth
    expression, _ = Grammar(problems.FAS_2D_BASIC, lo, hi, FAS=True).compile(golden["grammar"])
    ops = fas.lower(expression, template_smoother=golden["template_smoother"])
    solver = fas.FasSolver(lo, hi)
    history, ms = solver.solve(ops)
    u_ref, want = fo.solve_ops(ops, hi, lo)
    assert len(history) == len(want) and len(history) > 3
    assert np.all(np.abs(np.array(history) - np.array(want)) <= 1e-10 * np.array(want) + 1e-13 * want[0])
    assert close(solver.solution(), u_ref, 1e-10)
    solver.close()


@pytest.mark.gpu
def test_generator_mirror_evaluates_like_the_oracle():
    from evostencils_b200 import problems
    from evostencils_b200.grammar import Grammar
    from oracle import fas_oracle as fo
    gen = fas.B200ProgramGeneratorFAS(fct_smoother=None, min_level=4, max_level=7, device=0)
    golden = load_golden()["fas_inline_w_like"]
    expression, _ = Grammar(problems.FAS_2D_BASIC, 4, 7, FAS=True).compile(golden["grammar"])
    t, c, n = gen.generate_and_evaluate(expression, None, 4, 7, evaluation_samples=2)
    _, want = fo.solve_ops(gen.lower(expression), 7, 4)
    assert n == len(want) - 1 and t > 0
    # the last residuals sit at 1e-10 of the first, where exp() round-off shows: the n-th root agrees to 1e-3
    r0, rn = fas.printed_residuals(want)          # parse_output forms the factor from the printed (rounded) residuals
    assert abs(c - (rn / r0) ** (1.0 / n)) <= 1e-3 * c
    gen.close()


@pytest.mark.gpu
@pytest.mark.parametrize("pre,post", [(2, 1), (1, 1)])
def test_graph_replays_are_bitwise_the_direct_launches(pre, post):
    """The cycle captured as one CUDA graph per ping-pong state (V(2,1) with Jacobi smoothing alternates between
    two states, V(1,1) keeps one) against statement-by-statement launches: identical histories and fields."""
    lo, hi = 5, 8
    ops = fas.v_cycle_ops(hi, lo, pre, post)
    direct = fas.FasSolver(lo, hi, use_graph=False)
    want, _ = direct.solve(ops, max_iterations=9)
    u_want = direct.solution()
    direct.close()
    solver = fas.FasSolver(lo, hi)
    got, ms = solver.solve(ops, max_iterations=9)
    u_got = solver.solution()
    launches = solver.ws.launch_count()
    solver.close()
    assert got == want and len(got) == 10
    assert np.array_equal(u_got, u_want)
    assert ms > 0 and launches > 0


def test_convergence_factor_is_formed_from_the_printed_residuals():
    """parse_output (exastencils_FAS.py:370-393) reads what the `.exa4` prints: 4 significant digits, fewer for
    tiny residuals, nothing below 1e-12."""
    hist = [123.456789, 1.23456789, 4.56789e-10, 7.89123e-12, 3.3e-13]
    r0, rn = fas.printed_residuals(hist)
    assert r0 == 123.5
    assert rn == 8e-12                       # the last line printed: 1 significant digit; 3.3e-13 prints 'EFFECTIVELY ZERO'
    assert fas.printed_residuals([5.0, 7.89e-11])[1] == 7.9e-11 and fas.printed_residuals([5.0, 2.567e-10])[1] == 2.57e-10
    assert fas.printed_residuals([5.0, 1e-13]) == (5.0, 0.0)

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
    (kind, level of the field written, count or factor)."""
    import re
    out = []
    for line in text.splitlines():
        line = line.strip()
        m = re.match(r"Smoother@(\d+) \(", line)
        if m:
            level = int(m.group(1))
            if out and out[-1][0] == "smooth" and out[-1][1] == level:
                out[-1] = ("smooth", level, out[-1][2] + 1)
            else:
                out.append(("smooth", level, 1))
            continue
        m = re.match(r"CGS@(\d+) \(", line)
        if m:
            out.append(("cgs", int(m.group(1)), 1))
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
        assert line == "" or line == "}" or line.startswith(("Function", "loop over", "apply bc", "communicate",
                                                             "Solution@")), line
    return out


def test_cycle_statement_order_is_that_of_the_reference_emitter():
    """tests/golden/fas_golden.json: Layer-4 text emitted by the unmodified reference
    (`ProgramGeneratorFAS.generate_cycle_function`) for FAS V-cycles written in its FAS grammar."""
    import json
    import os
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "fas_golden.json")))
    assert len(golden) >= 4
    for name, case in golden.items():
        want = parse_layer4(case["layer4"])
        factors = {w for kind, _, w in want if kind == "prolong_add"}
        assert len(factors) == 1, name
        ops = fas.v_cycle_ops(case["max_level"], case["min_level"], case["pre"], case["post"], cgc_omega=factors.pop())
        got = []
        for op in ops:
            if op.kind == "smooth" and op.level == case["min_level"]:
                got.append(("cgs", op.level, 1))                    # CGS@coarsest = 30 sweeps of the smoother
            elif op.kind == "smooth":
                got.append(("smooth", op.level, op.sweeps))
            elif op.kind == "prolong_add":
                got.append(("prolong_add", op.level, op.omega))
            else:
                got.append((op.kind, op.level, 1))
        assert got == want, name


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

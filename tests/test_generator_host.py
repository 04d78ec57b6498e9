"""Host logic of B200ProgramGenerator.generate_and_evaluate on a mocked workspace (no GPU):
the k, 2k, 4k sweep lowers the individual again per wave number, per-individual engine errors give
the sentinel triple, a broken engine propagates."""
import numpy as np
import pytest

from evostencils_b200 import engine, lowering, problems
from evostencils_b200.generator import B200ProgramGenerator
from evostencils_b200.grammar import Grammar, v_cycle_string


class FakeWorkspace:
    """Records what the generator hands to the engine; `solve*` return a converging history."""

    def __init__(self, lo, hi, fail=None):
        self.min_level, self.max_level, self.n_fields, self.device = lo, hi, 1, 0
        self._operators = {}
        self.built = []
        self.outer = []
        self.fail = fail
        self.time_limit = None

    def set_operator(self, level, table):
        self._operators[level] = table

    def set_variable_coefficient(self, spec):
        self.variable = spec

    def set_boundary(self, *a):
        pass

    def set_coarse_solver(self, *a):
        pass

    def clear_programs(self):
        pass

    def set_time_limit(self, seconds):
        self.time_limit = seconds

    def build_cycle(self, ops, default_cycle=None):
        if self.fail is not None:
            raise self.fail
        self.built.append(list(ops))
        return len(self.built) - 1

    def set_outer_operator(self, entries):
        self.outer.append(entries)

    def upload(self, *a):
        pass

    def save_initial_guess(self):
        pass

    def restore_initial_guess(self):
        pass

    def solve_bicgstab(self, cycle, tol, max_its):
        return [1.0, 0.1, 1e-3, 1e-9], True, 2.0

    def solve(self, cycle, tol, max_its):
        return [1.0, 1e-5, 1e-13], True, 1.5

    def launch_count(self):
        return 7

    def close(self):
        pass


def make(problem, lo, hi, fail=None, **kw):
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, **kw)
    ws = FakeWorkspace(lo, hi, fail)
    gen._workspace, gen._workspace_key = ws, (lo, hi)
    return gen, ws


def test_wave_number_sweep_lowers_the_individual_for_every_k():
    """exastencils.py:518-532: k, 2k, 4k on the same grid.  The IR of an individual holds numbers, so
    operators AND local-solve matrices must follow the wave number (the sweep used to stop
    at 2k with 'the individual's operator differs from the problem's')."""
    problem, lo, hi, k = problems.HELMHOLTZ_2D, 3, 5, 5.0
    gen, ws = make(problem, lo, hi)
    expression, _ = Grammar(problem, lo, hi, parameters={"k": k}).compile(v_cycle_string(hi - lo, 2, 1, 10, "red_black"))
    t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1,
                                          global_variable_values={"k": k})
    assert (t, n) == (2.0, 2.0) and rho == pytest.approx(1e-9)
    assert [s[0]["k"] for s in gen.last_sweep] == [5.0, 10.0, 20.0]
    assert len(ws.built) == 3 and len(ws.outer) == 3
    h2 = 4.0 ** hi
    for i, kk in enumerate((5.0, 10.0, 20.0)):
        first = ws.built[i][0]
        assert isinstance(first, lowering.LocalSolve)
        assert first.matrix[0][0] == pytest.approx(4.0 * h2 - (1.0 + 0.5j) * kk * kk)      # shifted operator M
        centre = [v for o, v in ws.outer[i] if o == (0, 0)][0]
        assert centre == pytest.approx(4.0 * h2 - kk * kk)                                   # unshifted A
    # the level operators handed to the engine are the last wave number's
    table = lowering.operator_table(problem.system_operator(hi, 0, {"k": 20.0, "shift": 1.0 + 0.5j}))
    assert ws._operators[hi] == table
    # without k among the caller's values there is one run (exastencils.py:519-521)
    gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
    assert len(gen.last_sweep) == 1


def test_an_individual_built_for_another_wave_number_is_evaluated_at_the_requested_one():
    problem, lo, hi = problems.HELMHOLTZ_2D, 3, 4
    gen, ws = make(problem, lo, hi)
    expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 1, 1, 10, "single"))     # k = 80 baked in
    gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1, global_variable_values={"k": 7.0})
    assert ws.built[0][0].matrix[0][0] == pytest.approx(4.0 * 4.0 ** hi - (1.0 + 0.5j) * 49.0)


def test_per_individual_engine_errors_give_the_sentinel():
    problem, lo, hi = problems.POISSON_2D, 3, 5
    expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 1, 1, 21, "red_black"))
    for exc in (engine.EngineError("op kind 'richardson' is not executable"),
                engine.EngineError("mgengine error -1: bad op list", code=engine.MG_ERR_INVALID),
                engine.EngineError("mgengine error -5: solve exceeded the evaluation time limit", code=engine.MG_ERR_TIMEOUT)):
        gen, ws = make(problem, lo, hi, fail=exc)
        assert gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1) == (1e100,) * 3
        assert gen.last_error == str(exc)
    for exc in (engine.EngineError("mgengine error -2: cudaErrorIllegalAddress", code=engine.MG_ERR_CUDA),
                engine.EngineError("mgengine error -4: cudaMalloc failed", code=engine.MG_ERR_NOMEM),
                engine.EngineError("libmgengine.so not found", fatal=True)):
        gen, ws = make(problem, lo, hi, fail=exc)
        with pytest.raises(engine.EngineError):
            gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)


def test_the_time_limit_of_the_constructor_reaches_the_engine():
    problem, lo, hi = problems.POISSON_2D, 3, 5
    expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 1, 1, 21, "red_black"))
    gen, ws = make(problem, lo, hi, evaluation_timeout=42)
    t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=3)
    assert ws.time_limit == 42 and (t, n) == (1.5, 2.0)


def test_solo_timing_replays_the_cycle_under_the_exclusive_gate():
    from evostencils_b200 import population
    problem, lo, hi = problems.POISSON_2D, 3, 5
    expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 1, 1, 21, "red_black"))
    gen, ws = make(problem, lo, hi)
    replays = []
    ws.run_cycles = lambda cycle, n: replays.append(n) or 0.9
    gen.solo_timing = population.SoloGate()
    t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=2)
    assert replays == [3, 3] and t == pytest.approx(0.9 / 3 * 2)      # ms per replayed cycle x iterations

"""Host logic of bench.py that needs no GPU: the step of the population bench spreads its individuals
over the rank's generators (one host thread each) and counts cycles and finest-level launches."""
import os
import sys
import threading
import time

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, REPO)

import bench  # noqa: E402
from evostencils_b200 import lowering, problems  # noqa: E402
from evostencils_b200.grammar import Grammar, v_cycle_string  # noqa: E402


class FakeGenerator:
    """Stands in for B200ProgramGenerator: lowers for real, 'solves' in a fixed number of cycles."""

    def __init__(self):
        self.calls = []
        self.threads = set()
        self.last_residual_history = []
        self.last_program = None

    def generate_and_evaluate(self, expression, storages, lo, hi, solver, evaluation_samples=1):
        self.last_program = lowering.lower(expression, lo, hi)
        cycles = 3 + len(self.last_program.ops) % 4
        self.last_residual_history = [1.0] * (cycles + 1)
        self.threads.add(threading.get_ident())
        time.sleep(0.01)                        # long enough for the other threads to take jobs too
        self.calls.append(cycles)
        return (1.0, 0.1, float(cycles))


def test_evaluate_shard_spreads_the_step_over_the_generators():
    problem, lo, hi = problems.POISSON_2D, 7, 9
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    jobs = [v_cycle_string(2, pre, post, 21, part) for pre in (1, 2, 3) for post in (0, 1, 2) for part in ("single", "red_black")]
    for schedule in ("dynamic", "static"):
        for n_gens in (1, 3):
            gens = [FakeGenerator() for _ in range(n_gens)]
            counts = {}
            cycles, results = bench.evaluate_shard(gens, grammar, jobs, lo, hi, counts, schedule=schedule)
            assert len(results) == len(jobs)
            assert sum(len(g.calls) for g in gens) == len(jobs)                  # every individual exactly once
            assert cycles == sum(sum(g.calls) for g in gens)
            assert set(counts) <= {"jacobi", "rbgs", "residual", "residual_restrict", "prolong_add", "block_jacobi"}
            assert counts["prolong_add"] == cycles and counts["residual_restrict"] == cycles     # one of each per V-cycle
            if n_gens > 1:
                assert len(set().union(*[g.threads for g in gens])) == n_gens      # one host thread per generator


def test_finest_stage_counts_of_a_v_cycle():
    problem, lo, hi = problems.POISSON_2D, 7, 9
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    program = lowering.lower(grammar.compile(v_cycle_string(2, 2, 1, 21, "red_black"))[0], lo, hi)
    assert bench.finest_stage_counts(program, hi) == {"rbgs": 3, "residual_restrict": 1, "prolong_add": 1}
    program = lowering.lower(grammar.compile(v_cycle_string(2, 2, 2, 14, "single"))[0], lo, hi)
    assert bench.finest_stage_counts(program, hi) == {"jacobi": 4, "residual_restrict": 1, "prolong_add": 1}

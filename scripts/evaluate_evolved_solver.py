#!/usr/bin/env python
"""B200 counterpart of the reference's scripts/evaluate_evolved_solver.py: evaluate a persisted
individual (`hof_*/individual_*.txt`, the grammar string `str(individual)` the optimizer writes)
through `generate_and_evaluate` -- the path of
`Optimizer.generate_and_evaluate_program_from_grammar_representation` (optimization/program.py:904-929)
-- and print the reference's line.  With `--exa` the solver is also written as the ExaSlang cycle
function the reference would emit (`generate_cycle_function`, exastencils.py:318-336)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("individual", help="file holding the grammar string of one individual")
    ap.add_argument("--problem", default="2D_FD_Poisson_fromL2")
    ap.add_argument("--min-level", type=int, default=None)
    ap.add_argument("--max-level", type=int, default=None)
    ap.add_argument("--maximum-block-size", type=int, default=8, help="reference: maximum_block_size = 8")
    ap.add_argument("--evaluation-samples", type=int, default=20, help="reference: program.py:924")
    ap.add_argument("--engine-min-level", type=int, default=None,
                    help="let the engine's hierarchy go below the individual's coarsest level (DESIGN.md section 6)")
    ap.add_argument("--exa", default=None, help="write the ExaSlang cycle function to this file")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    from evostencils_b200 import problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar
    problem = problems.get_problem(a.problem)
    lo = problem.min_level if a.min_level is None else a.min_level
    hi = problem.max_level if a.max_level is None else a.max_level
    with open(a.individual) as fh:
        grammar_string = fh.read().strip()
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=a.maximum_block_size)
    expression, _ = grammar.compile(grammar_string)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=a.device,
                               engine_min_level=a.engine_min_level)
    if a.exa:
        with open(a.exa, "w") as fh:
            fh.write(gen.generate_cycle_function(expression, None, lo, hi, hi))
    gen.initialize_code_generation(lo, hi)
    solving_time, convergence_factor, number_of_iterations = gen.generate_and_evaluate(
        expression, None, lo, hi, "", evaluation_samples=a.evaluation_samples)
    gen.close()
    print(f"Solving Time: {solving_time}, "
          f"Convergence factor: {convergence_factor}, "
          f"Number of Iterations: {number_of_iterations}", flush=True)
    return solving_time, convergence_factor, number_of_iterations


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""B200 counterpart of the reference's scripts/evaluate_reference_solver.py (:5-47).

The reference compiles `Poisson/2D_FD_Poisson_fromL2` with ExaStencils -- i.e. the solver its `.exa3`
block describes (`generate solver ... with {...}`: V(numPre, numPost), coloured point smoother with
`damping`, CG on the coarsest level) -- runs the binary 20 times and prints the average solving time
and iteration count.  Here the same default cycle is lowered (`lowering.default_cycle_ops`) and solved
on the GPU engine; the two output lines are the reference's.  No java, make or g++."""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--problem", default="2D_FD_Poisson_fromL2", help="configuration name (reference: dir_name/problem_name)")
    ap.add_argument("--min-level", type=int, default=None, help="default: the problem's knowledge file value")
    ap.add_argument("--max-level", type=int, default=None)
    ap.add_argument("--nruns", type=int, default=20, help="reference: nruns = 20 (:15)")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    problem = problems.get_problem(a.problem)
    lo = problem.min_level if a.min_level is None else a.min_level
    hi = problem.max_level if a.max_level is None else a.max_level
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=a.device)
    gen.initialize_code_generation(lo, hi)
    ops = lowering.default_cycle_ops(problem, hi, lo)
    operators = {level: lowering.operator_table(problem.system_operator(level, 0)) for level in range(lo, hi + 1)}
    program = lowering.Program(ops, lo, hi, problem.n_fields, problem.dimension, operators)
    total_time = total_iterations = 0.0
    for _ in range(a.nruns):
        ms, rho, n = gen.evaluate_program(program, {}, evaluation_samples=1)
        total_time += ms
        total_iterations += n if n < 1e99 else 100          # reference :38-41
    gen.close()
    print(f"Average solving time: {total_time / a.nruns}", flush=True)
    print(f"Average number of iterations: {total_iterations / a.nruns}", flush=True)
    return total_time / a.nruns, total_iterations / a.nruns


if __name__ == "__main__":
    main()

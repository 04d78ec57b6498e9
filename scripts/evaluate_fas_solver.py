#!/usr/bin/env python
"""The reference's nonlinear example `example_problems/FAS_2D_Basic` on the GPU engine: FAS V(pre, post)
cycle with the damped Jacobi-Newton smoother of `FAS_2D_Basic.exa4`, solved until the `.exa4`'s
stopping rule holds; prints the two lines the generated binary prints (:185-186), which
`ProgramGeneratorFAS.generate_and_evaluate` parses (exastencils_FAS.py:369-426).  Levels default to
the `.knowledge` file's (minLevel 6, maxLevel 10).  (The CPU figure of the same solve is timed by
`bench.py --config fas`, not here: nothing under scripts/ touches oracle/.)"""
import argparse
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--min-level", type=int, default=6)
    ap.add_argument("--max-level", type=int, default=10)
    ap.add_argument("--pre", type=int, default=2)
    ap.add_argument("--post", type=int, default=2)
    ap.add_argument("--nruns", type=int, default=5)
    ap.add_argument("--individual", default=None,
                    help="file with the grammar string of an evolved FAS solver (FAS productions, depth = max - min level): "
                         "evaluated through B200ProgramGeneratorFAS.generate_and_evaluate instead of the V(pre, post) cycle")
    ap.add_argument("--template-smoother", action="store_true",
                    help="with --individual: every smoothing step calls the template's Smoother (fct_smoother given)")
    ap.add_argument("--device", type=int, default=0)
    a = ap.parse_args(argv)
    from evostencils_b200 import fas
    if a.individual:
        from evostencils_b200 import problems
        from evostencils_b200.grammar import Grammar
        with open(a.individual) as fh:
            string = fh.read().strip()
        expression, _ = Grammar(problems.FAS_2D_BASIC, a.min_level, a.max_level, FAS=True).compile(string)
        gen = fas.B200ProgramGeneratorFAS(fct_smoother="Smoother" if a.template_smoother else None,
                                          min_level=a.min_level, max_level=a.max_level, device=a.device)
        t, c, n = gen.generate_and_evaluate(expression, evaluation_samples=a.nruns)
        gen.close()
        print("total no of iterations", n)
        print("time to solution (in ms) is", t)
        print(f"convergence factor {c}")
        return n, t
    solver = fas.FasSolver(a.min_level, a.max_level, device=a.device)
    ops = fas.v_cycle_ops(a.max_level, a.min_level, a.pre, a.post)
    solver.solve(ops)                                   # warm-up
    best = None
    for _ in range(a.nruns):
        solver.reset()
        history, ms = solver.solve(ops)
        best = ms if best is None else min(best, ms)
    error = float(abs(solver.solution() - fas.exact_solution(a.max_level)).max())
    solver.close()
    print("total no of iterations", len(history) - 1)
    print("time to solution (in ms) is", best)
    print(f"residual reduction {history[-1] / history[0]:.3e}, max error against solFct {error:.3e}")
    return len(history) - 1, best


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""One elasticity solve (levels 4..8, V(2,1) collective RB-GS w = 1.25): ms per solve and per cycle."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar, v_cycle_string
    problem, lo, hi = problems.LINEAR_ELASTICITY_2D, 4, 8
    grammar = Grammar(problem, lo, hi)
    program = lowering.lower(grammar.compile(v_cycle_string(hi - lo, 2, 1, 23, "red_black"))[0], lo, hi)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.evaluate_program(program, {}, evaluation_samples=1)
    t, rho, n = gen.evaluate_program(program, {}, evaluation_samples=3)
    print(json.dumps({"ms_per_solve": t, "iterations": n, "ms_per_cycle": t / n}))
    gen.close()


if __name__ == "__main__":
    main()

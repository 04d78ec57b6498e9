#!/usr/bin/env python
"""One Helmholtz solve (levels 3..7, k = 80, V(2,1) RB-GS preconditioner inside BiCGStab), capped at a few outer
iterations: target of an ncu launch list (which kernels an iteration is made of and how long each takes)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar, v_cycle_string
    its = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    problem, lo, hi = problems.HELMHOLTZ_2D, 3, 7
    params = {"k": 80.0}
    grammar = Grammar(problem, lo, hi, parameters=params)
    program = lowering.lower(grammar.compile(v_cycle_string(hi - lo, 2, 1, 10, "red_black"))[0], lo, hi)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    t, rho, n = gen.evaluate_program(program, params, evaluation_samples=1, max_iterations=its)
    t, rho, n = gen.evaluate_program(program, params, evaluation_samples=1, max_iterations=its)
    print(json.dumps({"ms": t, "ms_per_iteration": t / its, "launches": gen.last_launch_count}))
    gen.close()


if __name__ == "__main__":
    main()

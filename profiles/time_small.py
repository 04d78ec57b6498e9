#!/usr/bin/env python
"""configs[0] (2D Poisson 513^2, levels 5..9, V(2,1) RB-GS): where a solve's time goes -- back-to-back graph
replays, the solve loop (graph launch + stream sync + host check per iteration), the first solve after the
cycle was built (capture + instantiation inside the call), and a per-op breakdown with CUDA events."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar, v_cycle_string
    problem, lo, hi = problems.POISSON_2D, 5, 9
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0)
    gen.initialize_code_generation(lo, hi)
    ws = gen._ensure_workspace(lo, hi)
    gen._configure(ws, {})
    us, fs = problem.initial_fields(hi)
    program = lowering.lower(Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))[0], lo, hi)
    out = {}
    for rep in range(2):
        ws.clear_programs()
        ws.upload("sol", hi, 0, us[0]); ws.upload("rhs", hi, 0, fs[0])
        cyc = ws.build_cycle(program.ops)
        hist, conv, ms = ws.solve(cyc, 1e-12, 100)
        out[f"first_solve_after_build_{rep}"] = {"ms": round(ms, 4), "iterations": len(hist) - 1}
        ws.upload("sol", hi, 0, us[0])
        hist, conv, ms = ws.solve(cyc, 1e-12, 100)
        out[f"second_solve_{rep}"] = {"ms": round(ms, 4), "iterations": len(hist) - 1}
    out["replay_ms_per_cycle"] = round(ws.run_cycles(cyc, 50) / 50, 4)
    per_op = ws.profile_ops(program.ops)
    names = ["smooth", "residual", "restrict", "residual+restrict", "zero", "prolong+add", "prolong", "cgs", "krylov"]
    out["per_op_ms"] = [[names[k], level, round(t, 4)] for k, level, _, t in per_op]
    out["per_op_sum_ms"] = round(sum(t for _, _, _, t in per_op), 4)
    print(json.dumps(out))
    gen.close()


if __name__ == "__main__":
    main()

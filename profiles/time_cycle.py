#!/usr/bin/env python
"""ms per captured-graph cycle of bench individuals at 4096^2 (CUDA events around 20 replays)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench


def main():
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar
    problem = problems.POISSON_2D
    hi, lo, elo = 12, 7, 2
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, engine_min_level=elo)
    gen.initialize_code_generation(lo, hi)
    ws = gen._ensure_workspace(lo, hi)
    gen._configure(ws, {})
    us, fs = bench.synthetic_inputs(problem, hi, pinned=False)
    ws.upload("sol", hi, 0, us[0]); ws.upload("rhs", hi, 0, fs[0])
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    pop = bench.load_population()[:8]
    default_cycle = lambda level: lowering.default_cycle_ops(problem, level, elo, {})
    out, solve_ms = [], []
    for string in pop:
        program = lowering.lower(grammar.compile(string)[0], lo, hi)
        ws.clear_programs()
        cyc = ws.build_cycle(program.ops, default_cycle)
        ws.run_cycles(cyc, 3)
        out.append(round(ws.run_cycles(cyc, 20) / 20, 4))
        ws.upload("sol", hi, 0, us[0])
        hist, conv, ms = ws.solve(cyc, 1e-12, 100)          # one graph launch + stream sync + host check per iteration
        solve_ms.append(round(ms / (len(hist) - 1), 4))
    print(json.dumps({"tail_level": os.environ.get("MG_TAIL_LEVEL", "default"), "ms_per_cycle": out,
                      "ms_per_cycle_in_solve": solve_ms}))
    gen.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Per-stage device time of one cycle of each bench individual (CUDA events after every op through
mg_profile_ops; no profiler).  Prints, per individual, the cycle time split by (stage, level) and
the total against the HBM-roofline time of the same op list (SURVEY.md 8d byte table)."""
import argparse
import collections
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

KIND = ["smooth", "residual", "restrict", "residual_restrict", "zero", "prolong_add", "prolong", "cgs"]
BYTES = {"smooth": 24.0, "residual": 24.0, "restrict": 10.0, "residual_restrict": 18.0, "zero": 8.0,
         "prolong_add": 18.0, "prolong": 10.0, "cgs": 0.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--individuals", type=int, default=8)
    ap.add_argument("--max-level", type=int, default=12)
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar
    problem = problems.POISSON_2D
    hi, lo, elo = a.max_level, a.max_level - 5, 2
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=0, engine_min_level=elo)
    gen.initialize_code_generation(lo, hi)
    ws = gen._ensure_workspace(lo, hi)
    gen._configure(ws, {})
    us, fs = bench.synthetic_inputs(problem, hi, pinned=False)
    ws.upload("sol", hi, 0, us[0])
    ws.upload("rhs", hi, 0, fs[0])
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    peak, _ = bench.measured_peaks()
    pop = bench.load_population()[: a.individuals]
    default_cycle = lambda level: lowering.default_cycle_ops(problem, level, elo, {})
    grand = collections.Counter()
    for idx, string in enumerate(pop):
        program = lowering.lower(grammar.compile(string)[0], lo, hi)
        ws.clear_programs()
        best = None
        for _ in range(a.reps):
            prof = ws.profile_ops(program.ops, default_cycle)
            if best is None or sum(p[3] for p in prof) < sum(p[3] for p in best):
                best = prof
        agg = collections.Counter()
        ideal = 0.0
        tiles = {}
        for kind, level, handle, ms in best:
            name = KIND[kind]
            level_of_work = level + 1 if name in ("restrict", "residual_restrict") else level
            key = (name, level_of_work) if level_of_work >= 8 else ("coarse<=7", 0)
            agg[key] += ms
            grand[key] += ms
            dof = float((2 ** level_of_work - 1) ** 2)
            ideal += BYTES[name] * dof / (peak * 1e9) * 1e3
        total = sum(agg.values())
        top = sorted(agg.items(), key=lambda kv: -kv[1])[:8]
        print(json.dumps({"individual": idx, "ops": len(best), "cycle_ms": round(total, 4),
                          "roofline_ms": round(ideal, 4), "frac": round(ideal / total, 3),
                          "top": [[f"{k[0]}@{k[1]}", round(v, 4)] for k, v in top]}))
    tot = sum(grand.values())
    print(json.dumps({"all": [[f"{k[0]}@{k[1]}", round(v, 4), round(v / tot, 3)]
                              for k, v in sorted(grand.items(), key=lambda kv: -kv[1])[:20]]}))
    gen.close()


if __name__ == "__main__":
    main()

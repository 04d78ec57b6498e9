#!/usr/bin/env python
"""The replicated part of a decomposed 513^3 cycle at 8 GPUs = a whole V(2,1) RB-GS cycle on 129^3 (levels 2..7):
graph-replayed time and a per-op breakdown (CUDA event after every op)."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=7)
    a = ap.parse_args()
    import bench_dd
    from evostencils_b200 import lowering, problems
    from evostencils_b200.engine import Workspace
    problem, lo, hi = problems.POISSON_3D, 2, a.level
    ws = Workspace(3, lo, hi, 1, "f64", 0)
    for level in range(lo, hi + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
    ops = bench_dd.cycle_ops(problem, hi, lo, 1)
    us, fs = problem.initial_fields(hi, {})
    ws.upload("rhs", hi, 0, fs[0]); ws.upload("sol", hi, 0, us[0])
    cyc = ws.build_cycle(ops)
    ws.run_cycles(cyc, 3)
    out = {"level": hi, "replay_ms_per_cycle": round(ws.run_cycles(cyc, 20) / 20, 4)}
    names = ["smooth", "residual", "restrict", "residual+restrict", "zero", "prolong+add", "prolong", "cgs", "krylov"]
    best = None
    for _ in range(3):
        prof = ws.profile_ops(ops)
        if best is None or sum(p[3] for p in prof) < sum(p[3] for p in best):
            best = prof
    out["per_op_ms"] = [[names[k], level, round(t, 4)] for k, level, _, t in best]
    out["per_op_sum_ms"] = round(sum(p[3] for p in best), 4)
    print(json.dumps(out))
    ws.close()


if __name__ == "__main__":
    main()

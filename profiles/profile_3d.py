#!/usr/bin/env python
"""Per-op device time of one V(2,1) RB-GS cycle of 3D Poisson (CUDA event after every op)."""
import argparse, collections, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench_dd

KIND = ["smooth", "residual", "restrict", "residual_restrict", "zero", "prolong_add", "prolong", "cgs"]
BYTES = {"smooth": 24.0, "residual": 24.0, "restrict": 9.0, "residual_restrict": 17.0, "zero": 8.0,
         "prolong_add": 17.0, "prolong": 9.0, "cgs": 0.0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--fused", type=int, default=1)
    a = ap.parse_args()
    from evostencils_b200 import lowering, problems
    from evostencils_b200.engine import Workspace
    problem = problems.POISSON_3D
    lo = 2
    ws = Workspace(3, lo, a.level, 1, "f64", 0)
    for level in range(lo, a.level + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver("CG", 1000, 1e-12)
    rhs, u0 = bench_dd.slab_inputs(problem, a.level)
    n = 2 ** a.level + 1
    ws.upload("rhs", a.level, 0, rhs[0](0, n))
    ws.upload("sol", a.level, 0, u0[0](0, n))
    ops = bench_dd.cycle_ops(problem, a.level, lo, 1)
    ws.profile_ops(ops, fused=bool(a.fused))
    prof = ws.profile_ops(ops, fused=bool(a.fused))
    agg = collections.Counter()
    ideal = 0.0
    for kind, level, handle, ms in prof:
        name = KIND[kind]
        lw = level + 1 if name in ("restrict", "residual_restrict") else level
        agg[(name, lw) if lw >= 7 else ("coarse<=6", 0)] += ms
        ideal += BYTES[name] * float((2 ** lw - 1) ** 3) / 6554.9e9 * 1e3
    total = sum(agg.values())
    print(json.dumps({"level": a.level, "cycle_ms": round(total, 3), "roofline_ms": round(ideal, 3),
                      "ops": [[f"{k[0]}@{k[1]}", round(v, 3)] for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:12]]}))
    ws.close()


if __name__ == "__main__":
    main()

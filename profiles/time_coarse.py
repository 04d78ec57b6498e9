#!/usr/bin/env python
"""Device time of the sub-cycle below a given level (default V(2,1) RB-GS cycle down to level 2),
with the single-CTA tail kernel and with one launch per op.  CUDA events, 50 repetitions."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--top", type=int, default=7)
    a = ap.parse_args()
    from evostencils_b200 import lowering, problems
    from evostencils_b200.engine import Workspace
    problem = problems.POISSON_2D
    lo, hi = 2, a.top + 1
    ws = Workspace(2, lo, hi, 1, "f64", 0)
    for level in range(lo, hi + 1):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_coarse_solver("CG", 1000, 1e-12)
    rng = np.random.default_rng(0)
    for level in range(lo, hi + 1):
        n = 2 ** level + 1
        f = np.zeros((n, n)); f[1:-1, 1:-1] = rng.standard_normal((n - 2, n - 2))
        ws.upload("rhs", level, 0, f)
        ws.upload("sol", level, 0, np.zeros((n, n)))
    out = {}
    for top in range(2, a.top + 1):
        ops = [lowering.Op("zero", top, dst="sol")] + lowering.fuse(lowering.default_cycle_ops(problem, top, lo))
        row = {}
        for name, tail in (("tail", True), ("launches", False)):
            ws.set_fast_kernels(True, tail=tail)
            ws.time_ops(ops, 5)
            row[name] = round(ws.time_ops(ops, 50) / 50 * 1e3, 2)
            cyc = ws.build_cycle(ops)
            ws.run_cycles(cyc, 5)
            row[name + "_graph"] = round(ws.run_cycles(cyc, 50) / 50 * 1e3, 2)      # + 2 launches of the level-(top+1) norm
        out[f"top{top}"] = row
    print(json.dumps({"us_per_subcycle": out, "tail_level": os.environ.get("MG_TAIL_LEVEL", "default")}))
    ws.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""3D prolongation + correction alone at 2^level+1 cubed: CUDA events on the engine stream, `--reps`
back-to-back launches after a warm-up.  Env MG_PROLONG3D_RY."""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    from evostencils_b200 import lowering, problems
    from evostencils_b200.engine import Workspace
    problem = problems.POISSON_3D
    ws = Workspace(3, a.level - 1, a.level, 1, "f64", 0)
    for level in (a.level - 1, a.level):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    n = 2 ** a.level + 1
    plane = np.random.default_rng(1).standard_normal((n, n))
    ws.upload("rhs", a.level, 0, np.broadcast_to(plane, (n, n, n)))
    ops = [lowering.Op("prolong_add", a.level, omega=1.0, src="sol", dst="sol")]
    ws.time_ops(ops, 3)
    ms = ws.time_ops(ops, a.reps) / a.reps
    dof = float((n - 2) ** 3)
    print(json.dumps({"level": a.level, "ms": round(ms, 4), "algorithmic_gbs": round(17 * dof / ms / 1e6, 1),
                      "env": {k: v for k, v in os.environ.items() if k.startswith("MG_")}}))
    ws.close()


if __name__ == "__main__":
    main()

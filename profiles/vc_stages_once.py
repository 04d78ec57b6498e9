#!/usr/bin/env python
"""The 2D variable-coefficient stage kernels alone at 4096^2: Jacobi sweep, red-black sweep (two colour passes),
residual -- event timing, and the ncu target (`-k regex:k_vcstar2d -c 4`)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    from evostencils_b200 import ir, lowering, problems
    from evostencils_b200.engine import Workspace
    problem, hi = problems.POISSON_2D_VARIABLE, 12
    ws = Workspace(2, hi - 1, hi, 1, "f64", 0)
    for level in (hi - 1, hi):
        ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
    ws.set_variable_coefficient(problem.variable_coefficient_spec())
    n = 2 ** hi + 1
    ws.upload("rhs", hi, 0, np.random.default_rng(1).standard_normal((n, n)))
    op = problem.system_operator(hi, 0)
    stages = {
        "jacobi": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, hi, 0.8, ir.Single),
        "rbgs": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, hi, 1.15, ir.RedBlack),
        "residual": [lowering.Op("residual", hi, dst="res")],
    }
    out = {}
    for name, ops in stages.items():
        if reps > 1:
            ws.time_ops(ops, 2)
        out[name] = round(ws.time_ops(ops, reps) / reps, 4)
    print(json.dumps({"ms": out, "env": {k: v for k, v in os.environ.items() if k.startswith("MG_")}}))
    ws.close()


if __name__ == "__main__":
    main()

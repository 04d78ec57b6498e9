import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from evostencils_b200 import lowering, problems
from evostencils_b200.engine import Workspace
problem = problems.POISSON_2D
lo, hi = 2, 7
ws = Workspace(2, lo, hi, 1, "f64", 0)
for level in range(lo, hi + 1):
    ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
ws.set_coarse_solver("CG", 1000, 1e-12)
rng = np.random.default_rng(0)
for level in range(lo, hi + 1):
    n = 2 ** level + 1
    f = np.zeros((n, n)); f[1:-1, 1:-1] = rng.standard_normal((n - 2, n - 2))
    ws.upload("rhs", level, 0, f); ws.upload("sol", level, 0, np.zeros((n, n)))
ops = [lowering.Op("zero", 6, dst="sol")] + lowering.fuse(lowering.default_cycle_ops(problem, 6, lo))
ws.exec_ops(ops)
ws.exec_ops(ops)
pass
ws.exec_ops(ops)
ws.close()

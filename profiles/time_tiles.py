#!/usr/bin/env python
"""Block-Jacobi tile kernel at 4096^2 for every tile shape with <= 4 unknowns (CUDA events, 20 launches)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench


def main():
    sys.argv = sys.argv[:1]
    args = bench.parse_args()
    from evostencils_b200 import ir, lowering
    peak, _ = bench.measured_peaks()
    gen, problem, lo, hi = bench.make_generator(args, 0, True)
    ws = gen._ensure_workspace(lo, hi)
    gen._configure(ws, {})
    us, fs = bench.synthetic_inputs(problem, hi, pinned=False)
    ws.upload("sol", hi, 0, us[0]); ws.upload("rhs", hi, 0, fs[0])
    op = problem.system_operator(hi, 0)
    dof = float((2 ** hi - 1) ** 2)
    out = {}
    for shape in ((1, 2), (2, 1), (1, 3), (3, 1), (2, 2), (1, 4), (4, 1)):
        ops = lowering.local_solve_statements(ir.generate_collective_block_jacobi(op, (shape,)), op, hi, 0.6, ir.Single)
        ws.time_ops(ops, 3)
        ws.flush_l2()
        ms = ws.time_ops(ops, 20) / 20
        out["%dx%d" % shape] = [round(ms * 1e3, 1), round(24.0 * dof / (ms * 1e-3) / 1e9 / peak, 3)]
    print(json.dumps({"us_and_fraction_of_copy_peak": out}))
    gen.close()


if __name__ == "__main__":
    main()

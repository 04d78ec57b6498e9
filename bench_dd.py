#!/usr/bin/env python
"""bench_dd.py -- BASELINE.json configs[2]: 3D Poisson 7-point, 2^L cells per dimension (default
L = 9: 512^3), fp64, V- or W-cycle with red-black GS (the `.exa3` solver block of
example_problems/Poisson/3D_FD_Poisson_fromL2.exa3: V(2,1), omega = 1.25, CG on level 2), domain
decomposed into z-slabs over the GPUs of one node with halo exchange over NCCL (strong scaling).

    python bench_dd.py --gpus 1 [--level 9] [--cycle V|W] [--cycles 5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench_dd.py --gpus N

Prints one JSON line: fine-grid DoF updates per second per cycle (whole job), ms per cycle (max over
ranks, device-synchronised), halo bytes per cycle, and at N = 1 the same cycle as a captured CUDA
graph of the single-domain engine for comparison."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def cycle_ops(problem, level, min_level, gamma):
    """V (gamma = 1) / W (gamma = 2) cycle of the `.exa3` solver block as an unfused op list."""
    from evostencils_b200 import ir, lowering
    s = problem.solver
    partitioning = ir.RedBlack if s.coloring == "red_black" else ir.Single
    ops = []

    def smooth(lv):
        operator = problem.system_operator(lv, 0)
        return lowering.local_solve_statements(ir.ElementwiseDiagonal(operator), operator, lv, s.damping, partitioning)

    def visit(lv):
        if lv == min_level:
            ops.append(lowering.Op("cgs", lv, src="rhs", dst="sol"))
            return
        for _ in range(s.num_pre):
            ops.extend(smooth(lv))
        ops.append(lowering.Op("residual", lv, dst="res"))
        ops.append(lowering.Op("restrict", lv - 1, src="res", dst="rhs"))
        ops.append(lowering.Op("zero", lv - 1, dst="sol"))
        for _ in range(gamma if lv - 1 > min_level else 1):
            visit(lv - 1)
        ops.append(lowering.Op("prolong_add", lv, omega=1.0, src="sol", dst="sol"))
        for _ in range(s.num_post):
            ops.extend(smooth(lv))

    visit(level)
    return ops


def slab_inputs(problem, level):
    """u0 = Dirichlet data of the example problem, interior 0; f ~ N(0,1) per plane (seed 1234 + z),
    so that every decomposition sees the same global field."""
    n = 2 ** level + 1
    x = np.linspace(0.0, 1.0, n)

    def rhs(z_lo, z_hi):
        out = np.zeros((z_hi - z_lo, n, n))
        for k, z in enumerate(range(z_lo, z_hi)):
            if 1 <= z <= n - 2:
                out[k, 1:-1, 1:-1] = np.random.default_rng(1234 + z).standard_normal((n - 2, n - 2))
        return out

    def u0(z_lo, z_hi):
        zz = x[z_lo:z_hi][:, None, None]
        full = x[None, None, :] ** 2 - 0.5 * x[None, :, None] ** 2 - 0.5 * zz ** 2 + np.zeros((z_hi - z_lo, n, n))
        out = np.zeros_like(full)
        out[:, 0, :], out[:, -1, :], out[:, :, 0], out[:, :, -1] = full[:, 0, :], full[:, -1, :], full[:, :, 0], full[:, :, -1]
        for k, z in enumerate(range(z_lo, z_hi)):
            if z == 0 or z == n - 1:
                out[k] = full[k]
        return out

    return [rhs], [u0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--level", type=int, default=9)
    ap.add_argument("--min-level", type=int, default=2)
    ap.add_argument("--cycle", default="V", choices=["V", "W"])
    ap.add_argument("--cycles", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--dd-min-level", type=int, default=None)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from evostencils_b200 import domain, lowering, problems
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    problem = problems.POISSON_3D
    ops = cycle_ops(problem, a.level, a.min_level, 1 if a.cycle == "V" else 2)
    comm = domain.SlabComm.from_environment()
    solver = domain.DomainDecomposedSolver(problem, a.min_level, a.level, comm, dd_min_level=a.dd_min_level,
                                           device_of=lambda r: local)
    rhs, u0 = slab_inputs(problem, a.level)
    solver.set_fields(rhs, u0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    r0 = solver.residual_norm()
    for _ in range(a.warmup):
        solver.run_ops(ops)
    comm.bytes_sent = 0
    solver.halo_exchanges = 0
    barrier()
    t0 = time.perf_counter()
    for _ in range(a.cycles):
        solver.run_ops(ops)
    barrier()
    ms = (time.perf_counter() - t0) * 1e3
    r1 = solver.residual_norm()
    stats = torch.tensor([ms, float(comm.bytes_sent)], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = stats.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = stats.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        ms, sent = mx[0].item(), sm[1].item()
    else:
        sent = stats[1].item()
    dof = float((2 ** a.level - 1) ** 3)
    line = {"metric": "mg_cycle_dof_updates_per_s", "value": dof * a.cycles / (ms * 1e-3), "unit": "DoF/s",
            "n_gpus": world, "scaling": "strong", "dtype": "f64", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[2]: 3D Poisson 7-point %d^3 fp64, %s(2,1)-cycle red-black GS "
                                   "omega=1.25, CG on level %d, z-slab domain decomposition" % (2 ** a.level, a.cycle, a.min_level),
                       "dd_min_level": solver.dd_min_level, "splits": solver.splits},
            "ms_per_cycle": ms / a.cycles, "cycles": a.cycles, "warmup": a.warmup,
            "halo_bytes_per_cycle": sent / a.cycles, "halo_exchanges_per_cycle": solver.halo_exchanges / a.cycles,
            "residual_reduction_over_run": r1 / r0,
            "timing": "host clock between device-synchronised barriers, max over ranks (op-by-op driver, no graph)"}
    solver.close()
    if world == 1:
        # the same cycle on the single-domain engine as one captured CUDA graph
        from evostencils_b200.engine import Workspace
        ws = Workspace(3, a.min_level, a.level, 1, "f64", local)
        for level in range(a.min_level, a.level + 1):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
        n = 2 ** a.level + 1
        ws.upload("rhs", a.level, 0, rhs[0](0, n))
        ws.upload("sol", a.level, 0, u0[0](0, n))
        cyc = ws.build_cycle(ops)
        ws.run_cycles(cyc, a.warmup)
        g_ms = ws.run_cycles(cyc, a.cycles) / a.cycles
        ws.close()
        line["single_domain_graph"] = {"ms_per_cycle": g_ms, "value": dof / (g_ms * 1e-3)}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- BASELINE.json configs[1]: fitness evaluation of an evolved-cycle population on
2D Poisson 4096^2 (fp64), sharded over the GPUs of one node (one process per GPU).

Step  = the first N x `--individuals-per-gpu` individuals of the stored 256-individual population
        (weak scaling; 32 per GPU = the whole population on 8 GPUs) are evaluated once through
        B200ProgramGenerator.generate_and_evaluate; ranks take individuals from a shared counter and
        evaluate `--streams` (default 2) of them at a time, each on its own workspace and CUDA stream.
value = fine-grid DoF updates per second = sum over individuals of (interior DoF x multigrid
        cycles executed) / step time, whole job, inputs resident in HBM.
e2e   = same metric with the RHS and the initial guess copied from pinned host memory for every
        evaluation and the residual history read back (the reference-facing call).
The line also carries `roofline` (dominant kernel, timed live with CUDA events), `cpu_baseline`
(C/OpenMP oracle port on the host cores, bounded sample), `evals_per_s`, `clocks`, `gpu_launches`.
`--impl reference` times the CPU port of the reference path instead (rank 0 only).
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "mg_cycle_dof_updates_per_s"
UNIT = "DoF/s"
POPULATION = os.path.join(ROOT, "evostencils_b200", "data", "population_2d_poisson_l7_12.txt")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--individuals-per-gpu", type=int, default=32,
                    help="weak scaling: a step evaluates the first N x this many individuals (256 / 8 GPUs = 32)")
    ap.add_argument("--schedule", default="dynamic", choices=["dynamic", "static"],
                    help="dynamic: ranks take the next individual of the step from a shared counter "
                         "(population.WorkQueue); static: the reference's i %% nprocs striding")
    ap.add_argument("--streams", type=int, default=2,
                    help="engine workspaces (= CUDA streams, one host thread each) per GPU: individuals of a rank are "
                         "evaluated concurrently, the coarse levels of one under the fine levels of another")
    ap.add_argument("--max-level", type=int, default=12)
    ap.add_argument("--grammar-depth", type=int, default=5)
    ap.add_argument("--engine-min-level", type=int, default=2)
    ap.add_argument("--cpu-sample-cycles", type=int, default=40,
                    help="CPU legs: cycles per sample (40 cycles = ~25 core-seconds on 16 cores)")
    ap.add_argument("--generic-kernels", action="store_true", help="disable the specialised kernels")
    ap.add_argument("--no-prefetch", action="store_true",
                    help="e2e leg: upload the inputs synchronously instead of one evaluation ahead on a copy stream")
    ap.add_argument("--stages-only", action="store_true",
                    help="time the finest-level stage kernels alone and print their rooflines (ncu target)")
    ap.add_argument("--stage-reps", type=int, default=20)
    return ap.parse_args()


def load_population():
    with open(POPULATION) as fh:
        return [line.strip() for line in fh if line.strip() and not line.startswith("#")]


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._thread = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
            except Exception:
                pass
            self._stop.wait(0.5)

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[3 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(sm)}


def synthetic_inputs(problem, level, pinned=True):
    """SURVEY.md 8d throughput inputs: u0 = 0, f ~ N(0,1) on interior nodes, boundary 0."""
    n = 2 ** level + 1
    rng = np.random.default_rng(1234)

    def host(shape):
        if pinned:
            import torch
            return torch.empty(shape, dtype=torch.float64, pin_memory=torch.cuda.is_available()).numpy()
        return np.empty(shape)

    f = host((n, n))
    f[...] = 0
    f[1:-1, 1:-1] = rng.standard_normal((n - 2, n - 2))
    u = host((n, n))
    u[...] = 0
    return [u], [f]


def captured_traffic(stage):
    """DRAM bytes per launch of the stage kernel from the committed `ncu --set full` capture
    (profiles/r1_traffic.json; read + write), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_traffic.json")) as f:
            return json.load(f)["stages"][stage]["traffic_bytes"]
    except (OSError, KeyError, ValueError):
        return None


def algorithmic_bytes_per_dof(stage, dim=2):
    """SURVEY.md 8d table (fp64)."""
    return {"jacobi": 24.0, "rbgs": 24.0, "block_jacobi": 24.0, "residual": 24.0, "residual_restrict": 16.0 + 8.0 / 2 ** dim,
            "prolong_add": 16.0 + 8.0 / 2 ** dim, "residual_norm": 16.0}[stage]


def stage_rooflines(gen, problem, level, peak_gbs, reps=20, warm=3):
    """Time every finest-level stage kernel alone (CUDA events on the engine stream, arrays of
    134 MB each > the 126 MB L2, 20 back-to-back launches after 3 warm-up launches)."""
    from evostencils_b200 import ir, lowering
    ws = gen._workspace
    op = problem.system_operator(level, 0)
    dof = float((2 ** level - 1) ** problem.dimension)
    stages = {
        "jacobi": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, level, 0.8, ir.Single),
        "rbgs": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, level, 1.15, ir.RedBlack),
        # block smoothers come in many tile shapes; 2x2 tiles stand for them in the share estimate
        "block_jacobi": lowering.local_solve_statements(ir.generate_collective_block_jacobi(op, ((2, 2),)), op, level,
                                                        0.6, ir.Single),
        "residual": [lowering.Op("residual", level, dst="res")],
        "residual_restrict": [lowering.Op("residual_restrict", level - 1, src="sol", dst="rhs")],
        "prolong_add": [lowering.Op("prolong_add", level, omega=1.0, src="sol", dst="sol")],
    }
    out = {}
    for name, ops in stages.items():
        ws.time_ops(ops, warm)
        ws.flush_l2()
        ms = ws.time_ops(ops, reps) / reps
        bytes_per_launch = algorithmic_bytes_per_dof(name, problem.dimension) * dof
        gbs = bytes_per_launch / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "achieved_gbs": gbs, "frac": gbs / peak_gbs, "bytes": bytes_per_launch}
    return out


def make_generator(args, device, resident):
    from evostencils_b200 import problems
    from evostencils_b200.generator import B200ProgramGenerator
    problem = problems.POISSON_2D
    hi = args.max_level
    lo = hi - args.grammar_depth
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=device,
                               engine_min_level=args.engine_min_level, resident_inputs=resident,
                               prefetch_inputs=not resident and not args.no_prefetch)
    gen.initialize_code_generation(lo, hi)
    if args.generic_kernels:
        gen._workspace.set_fast_kernels(False)
    return gen, problem, lo, hi


def finest_stage_counts(program, hi):
    """Finest-level launches of one cycle of `program`, by stage kernel."""
    from evostencils_b200 import lowering
    counts = {}
    for op in lowering.fuse(program.ops):
        if isinstance(op, lowering.LocalSolve):
            if op.level != hi:
                continue
            if any(p != 1 for p in op.period):
                name = "block_jacobi"
            else:
                name = "rbgs" if op.colored else "jacobi"
        elif op.kind in ("residual", "prolong_add") and op.level == hi:
            name = op.kind
        elif op.kind == "residual_restrict" and op.level == hi - 1:
            name = op.kind
        else:
            continue
        counts[name] = counts.get(name, 0) + 1
    return counts


def evaluate_shard(gens, grammar, jobs, lo, hi, stage_counts=None, schedule="dynamic"):
    """One step: this rank's part of the step's individuals `jobs` (self-scheduled over the ranks
    through population.WorkQueue, or strided i % world), spread over the rank's generators (one host
    thread, workspace and stream each).  -> (cycles executed, results)."""
    from evostencils_b200 import population
    if schedule == "dynamic":
        mine = population.WorkQueue(len(jobs))
    else:
        rank, world, _ = population._world()
        mine = population.shard_indices(len(jobs), rank, world)
    cycles = [0] * len(gens)
    lock = threading.Lock()

    def evaluator(slot, gen):
        def evaluate(string):
            expression, _ = grammar.compile(string)
            t, rho, n = gen.generate_and_evaluate(expression, None, lo, hi, "", evaluation_samples=1)
            c = max(0, len(gen.last_residual_history) - 1)
            cycles[slot] += c
            if stage_counts is not None and gen.last_program is not None:
                with lock:
                    for name, k in finest_stage_counts(gen.last_program, hi).items():
                        stage_counts[name] = stage_counts.get(name, 0) + k * c
            return (t, rho, n)
        return evaluate

    local = population.run_concurrently([evaluator(i, g) for i, g in enumerate(gens)], mine, jobs)
    return sum(cycles), [local[k] for k in sorted(local)]


def run_b200(args):
    import torch
    import torch.distributed as dist
    from evostencils_b200 import lowering
    from evostencils_b200.grammar import Grammar
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    population = load_population()
    p = args.individuals_per_gpu
    jobs = [population[i % len(population)] for i in range(world * p)]
    peak, peak_src = measured_peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_phase(resident):
        made = [make_generator(args, local, resident) for _ in range(max(1, args.streams))]
        gens = [m[0] for m in made]
        gen, problem, lo, hi = made[0]
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
        # longest-first by the static cost of one cycle (the number of cycles is not known beforehand)
        cost = [lowering.estimate_cycle_bytes(lowering.lower(grammar.compile(s)[0], lo, hi).ops, problem.dimension)
                for s in jobs]
        ordered = [jobs[i] for i in sorted(range(len(jobs)), key=lambda i: (-cost[i], i))]
        us, fs = synthetic_inputs(problem, hi)
        for g in gens:
            g.set_inputs(rhs=fs, u0=us)
        for _ in range(args.warmup):
            evaluate_shard(gens, grammar, ordered, lo, hi, schedule=args.schedule)
        wss = [g._workspace for g in gens]
        for g in gens:
            g.h2d_bytes = g.d2h_bytes = 0
        launches0 = sum(w.launch_count() for w in wss)
        barrier()
        for w in wss:
            w.mark(0)
        t0 = time.perf_counter()
        cycles = 0
        results = None
        counts = {}
        for _ in range(args.steps):
            c, results = evaluate_shard(gens, grammar, ordered, lo, hi, counts, schedule=args.schedule)
            cycles += c
        for w in wss:
            w.mark(1)
        for w in wss:
            w.synchronize()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        dev_ms = max(w.elapsed() for w in wss)          # every stream is marked at the common start and end
        launches = sum(w.launch_count() for w in wss) - launches0
        stats = torch.tensor([dev_ms, wall_ms, float(cycles), float(launches), float(sum(g.h2d_bytes for g in gens)),
                              float(sum(g.d2h_bytes for g in gens))], dtype=torch.float64, device="cuda")
        if world > 1:
            mx = stats.clone()
            dist.all_reduce(mx, op=dist.ReduceOp.MAX)
            sm = stats.clone()
            dist.all_reduce(sm, op=dist.ReduceOp.SUM)
            dev_ms, wall_ms = mx[0].item(), mx[1].item()
            cycles, launches = sm[2].item(), sm[3].item()
            h2d, d2h = sm[4].item(), sm[5].item()
        else:
            h2d, d2h = stats[4].item(), stats[5].item()
        return gens, problem, lo, hi, dev_ms, wall_ms, cycles, launches, h2d, d2h, results, counts

    # clocks are sampled for rank 0's GPU only: one nvidia-smi process per rank every 0.2 s takes the
    # driver's global locks often enough on an 8-GPU box to disturb the other ranks' launches
    with (ClockSampler(local) if rank == 0 else contextlib.nullcontext()) as clocks:
        gens, problem, lo, hi, dev_ms, wall_ms, cycles, launches, _, _, results, counts = timed_phase(resident=True)
        stages = stage_rooflines(gens[0], problem, hi, peak) if rank == 0 else None
        for g in gens:
            g.close()
        gens2, _, _, _, dev2, wall2, cycles2, launches2, h2d, d2h, _, _ = timed_phase(resident=False)
        for g in gens2:
            g.close()
    dof = float((2 ** hi - 1) ** 2)
    value = dof * cycles / (dev_ms * 1e-3)
    e2e = dof * cycles2 / (max(dev2, wall2) * 1e-3)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    converged = sum(1 for r in results if r[2] < 1e99)
    # share of rank 0's timed region spent in each finest-level stage kernel: launches counted from
    # the op lists x the stage's measured launch time; the largest is the dominant kernel
    share = {name: {"launches": k, "est_ms": k * stages[name]["ms"] if name in stages else None,
                    "share": k * stages[name]["ms"] / dev_ms if name in stages else None}
             for name, k in sorted(counts.items())}
    dominant = max((n for n in share if share[n]["share"] is not None), key=lambda n: share[n]["share"], default="rbgs")
    cpu = cpu_baseline(args, sample_only=True)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "BASELINE.json configs[1]: 2D Poisson 5-point 4096^2 fp64, fitness evaluation of "
                               "evolved-cycle individuals (grammar levels %d..%d, default V(2,1) RB-GS cycle + CG "
                               "below, engine levels %d..%d)" % (lo, hi, args.engine_min_level, hi),
                   "individuals_per_gpu": p, "individuals_per_step": world * p, "schedule": args.schedule, "streams_per_gpu": max(1, args.streams), "population": os.path.relpath(POPULATION, ROOT),
                   "evaluation_samples": 1, "target_reduction": 1e-12, "max_iterations": 100,
                   "inputs": "u0=0, f~N(0,1) seed 1234 (SURVEY.md 8d)",
                   "l2_policy": "every field array is 134 MB > 126 MB L2; stage timings additionally flush L2",
                   "kernels": "generic" if args.generic_kernels else "specialised+generic"},
        "evals_per_s": world * p * args.steps / (dev_ms * 1e-3),
        "cycles_per_step": cycles / args.steps, "converged_in_last_step_rank0": converged,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": h2d / args.steps,
                "d2h_bytes_per_step": d2h / args.steps, "ms_per_step": max(dev2, wall2) / args.steps,
                "input_copy": "synchronous before each solve" if args.no_prefetch else
                              "pinned host -> device staging on a copy stream, one evaluation ahead of the solve"},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": stages[dominant]["achieved_gbs"], "peak": peak,
                     "unit": "GB/s", "frac": stages[dominant]["frac"], "traffic": captured_traffic(dominant), "peak_source": peak_src,
                     "traffic_source": "profiles/r1_traffic.json (ncu --set full, same kernel and size)",
                     "algorithmic_bytes_per_launch": stages[dominant]["bytes"],
                     "ms_per_launch": stages[dominant]["ms"]},
        "stages": stages, "finest_level_share_rank0": share,
        "cpu_baseline": cpu,
        "clocks": clocks.summary(),
    }
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_stages(args):
    """`--stages-only`: the five finest-level stage kernels alone (used as the ncu target)."""
    import torch
    torch.cuda.set_device(0)
    peak, peak_src = measured_peaks()
    gen, problem, lo, hi = make_generator(args, 0, True)
    ws = gen._ensure_workspace(lo, hi)
    gen._configure(ws, {})
    us, fs = synthetic_inputs(problem, hi, pinned=False)
    ws.upload("sol", hi, 0, us[0])
    ws.upload("rhs", hi, 0, fs[0])
    stages = stage_rooflines(gen, problem, hi, peak, reps=args.stage_reps, warm=min(3, args.stage_reps))
    gen.close()
    print(json.dumps({"stages": stages, "peak": peak, "peak_source": peak_src, "level": hi, "reps": args.stage_reps}))


class CpuPort:
    """C/OpenMP oracle port on the host cores (all of them): cycles of the reference V(2,1) RB-GS
    individual (population[0]) at the bench size, one loop nest per unfused statement -- the
    granularity of the code ExaStencils generates.  Set up once, timed per call of `run`."""

    def __init__(self, args):
        from evostencils_b200 import engine, lowering, problems
        from evostencils_b200.grammar import Grammar
        oracle_dir = os.path.join(ROOT, "oracle")
        lib_path = os.path.join(oracle_dir, "_build", "libmgoracle.so")
        try:
            subprocess.run(["make", "-s", "-B", "-C", oracle_dir, "MARCH=native"], check=True, capture_output=True, timeout=120)
        except Exception:
            if not os.path.exists(lib_path):
                raise
        self.cores = len(os.sched_getaffinity(0))
        os.environ.setdefault("OMP_NUM_THREADS", str(self.cores))
        clib = engine.load_library(lib_path)
        problem = problems.POISSON_2D
        self.hi = hi = args.max_level
        lo = hi - args.grammar_depth
        elo = args.engine_min_level
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
        program = lowering.lower(grammar.compile(load_population()[0])[0], lo, hi)
        self.ws = ws = engine.Workspace(2, elo, hi, 1, "f64", 0, lib=clib)
        for level in range(elo, hi + 1):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.set_coarse_solver("CG", 1000, 1e-12)
        us, fs = synthetic_inputs(problem, hi, pinned=False)
        ws.upload("sol", hi, 0, us[0])
        ws.upload("rhs", hi, 0, fs[0])
        self.cycle = ws.build_cycle(program.ops, lambda level: lowering.default_cycle_ops(problem, level, elo), fused=False)
        ws.run_cycles(self.cycle, 1)                      # warm-up (page faults)

    def run(self, n):
        ms = self.ws.run_cycles(self.cycle, n)
        dof = float((2 ** self.hi - 1) ** 2)
        return {"value": dof * n / (ms * 1e-3), "unit": UNIT, "cores": self.cores, "kind": "port",
                "sample": "%d cycles (+ residual norm) of population[0] = V(2,1) RB-GS w=1.15 at %d^2, "
                          "C/OpenMP oracle port, unfused statements" % (n, 2 ** self.hi),
                "ms_per_cycle": ms / n}

    def close(self):
        self.ws.close()


def cpu_baseline(args, sample_only=False):
    port = CpuPort(args)
    try:
        return port.run(args.cpu_sample_cycles)
    finally:
        port.close()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    port = CpuPort(args)
    for _ in range(args.warmup):
        port.run(args.cpu_sample_cycles)
    t0 = time.perf_counter()
    vals = [port.run(args.cpu_sample_cycles) for _ in range(args.steps)]
    wall = (time.perf_counter() - t0) * 1e3
    port.close()
    value = float(np.mean([v["value"] for v in vals]))
    cpu = dict(vals[-1])
    cpu["value"] = value
    hi = args.max_level
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "BASELINE.json configs[1]: 2D Poisson 5-point %d^2 fp64; CPU arm = bounded sample "
                                   "(%d cycles of the V(2,1) RB-GS reference individual per step)" % (2 ** hi, args.cpu_sample_cycles),
                       "note": "the reference's own ExaStencils-generated C++ cannot be built here (no java); "
                               "this is the C/OpenMP oracle port (oracle/mg_oracle_c.c)"},
            "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.stages_only:
        run_stages(a)
    else:
        run_b200(a)

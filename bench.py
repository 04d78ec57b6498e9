#!/usr/bin/env python
"""bench.py -- BASELINE.json configs[1]: fitness evaluation of an evolved-cycle population on
2D Poisson 4096^2 (fp64), sharded over the GPUs of one node (one process per GPU).

Step  = the first N x `--individuals-per-gpu` individuals of the stored 256-individual population
        (weak scaling; 32 per GPU = the whole population on 8 GPUs) are evaluated once through
        B200ProgramGenerator.generate_and_evaluate; ranks take individuals from a shared counter and
        evaluate `--streams` (default 3) of them at a time, each on its own workspace and CUDA stream.
        `--strong`: every step evaluates all 256 individuals whatever N is.
value = fine-grid DoF updates per second = sum over individuals of (interior DoF x multigrid
        cycles executed) / step time, whole job, inputs resident in HBM.
e2e   = same metric with the RHS and the initial guess copied from pinned host memory for every
        evaluation and the residual history read back (the reference-facing call).
The line also carries
  roofline      dominant kernel, timed live with CUDA events, + `step`: algorithmic bytes of the
                whole timed region (sum over the individuals' op lists x cycles) / time / peak;
  cpu_baseline  C/OpenMP oracle port on the host cores: one cycle of each of the step's first 32
                individuals, with all host threads and with the reference's own 4
                (example_problems/lib/parallelization_pureOmp.knowledge:3);
  strong_scaling  one pass over all 256 individuals (the fixed population of configs[1]);
  dd3d          BASELINE.json configs[2]: one 513^3 V(2,1) RB-GS cycle, z-slabs over the N GPUs with
                halo exchange (N = 1: the single-domain engine);
  small_configs configs[0], [3], [4] (513^2 reference solver, Helmholtz 3..7 k = 80, elasticity
                4..8): ms per solve and evaluations/s on the GPU, the CPU figure beside each;
  variable_coefficient  2D variable-coefficient Poisson at the same size: stage rooflines and one solve;
  evals_per_s, clocks, gpu_launches.
`--impl reference` times the CPU port of the reference path instead (rank 0 only), on the same
`config`, one cycle of each of the same individuals per step.
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
    ap.add_argument("--streams", type=int, default=3,
                    help="engine workspaces (= CUDA streams, one host thread each) per GPU: individuals of a rank are "
                         "evaluated concurrently, the coarse levels of one under the fine levels of another")
    ap.add_argument("--max-level", type=int, default=12)
    ap.add_argument("--grammar-depth", type=int, default=5)
    ap.add_argument("--engine-min-level", type=int, default=2)
    ap.add_argument("--cpu-sample-individuals", type=int, default=32,
                    help="CPU legs: one cycle of each of the step's first this-many individuals per sample")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: every step evaluates all 256 individuals of the population at any N")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the strong-scaling pass, the 513^3 domain-decomposition leg and the small configurations")
    ap.add_argument("--dd-only", action="store_true", help="run the 513^3 domain-decomposition leg alone and print its record")
    ap.add_argument("--dd-level", type=int, default=9, help="finest level of the 3D leg (9 = 513^3)")
    ap.add_argument("--dd-cycles", type=int, default=5)
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


def cycle_bytes_of(program, problem, engine_min_level):
    """Algorithmic HBM bytes of one outer iteration of `program`: its op list, the
    default cycle the engine runs below the grammar's coarsest level, + 16 B/DoF for the fused
    residual norm that ends every iteration."""
    from evostencils_b200 import lowering
    default = lambda level: (lowering.default_cycle_ops(problem, level, engine_min_level)
                             if level > engine_min_level else None)
    dof = float((2 ** program.max_level - 1) ** problem.dimension)
    return lowering.estimate_cycle_bytes(program.ops, problem.dimension, default_cycle=default) + 16.0 * dof


def evaluate_shard(gens, grammar, jobs, lo, hi, stage_counts=None, schedule="dynamic", accounting=None):
    """One step: this rank's part of the step's individuals `jobs` (self-scheduled over the ranks
    through population.WorkQueue, or strided i % world), spread over the rank's generators (one host
    thread, workspace and stream each).  -> (cycles executed, results).
    `accounting` = {"bytes": 0.0, "cycle_bytes": callable(program)}: algorithmic bytes of the
    evaluations are added up (cycles x bytes of one iteration + the initial residual norm)."""
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
            if gen.last_program is not None and (stage_counts is not None or accounting is not None):
                with lock:
                    if stage_counts is not None:
                        for name, k in finest_stage_counts(gen.last_program, hi).items():
                            stage_counts[name] = stage_counts.get(name, 0) + k * c
                    if accounting is not None:
                        dof = float((2 ** hi - 1) ** gen.last_program.dimension)
                        accounting["bytes"] += c * accounting["cycle_bytes"](gen.last_program) + 16.0 * dof
            return (t, rho, n)
        return evaluate

    local = population.run_concurrently([evaluator(i, g) for i, g in enumerate(gens)], mine, jobs)
    return sum(cycles), [local[k] for k in sorted(local)]


def workload_config(args, world):
    """`config` of the JSON line: the same object for the GPU arm and the `--impl reference` arm."""
    hi = args.max_level
    lo = hi - args.grammar_depth
    p = 256 if args.strong else args.individuals_per_gpu * world
    return {"workload": "BASELINE.json configs[1]: 2D Poisson 5-point 4096^2 fp64, fitness evaluation of "
                        "evolved-cycle individuals (grammar levels %d..%d, default V(2,1) RB-GS cycle + CG "
                        "below, engine levels %d..%d)" % (lo, hi, args.engine_min_level, hi),
            "individuals_per_gpu": None if args.strong else args.individuals_per_gpu, "individuals_per_step": p,
            "population": os.path.relpath(POPULATION, ROOT),
            "evaluation_samples": 1, "target_reduction": 1e-12, "max_iterations": 100,
            "inputs": "u0=0, f~N(0,1) seed 1234 (SURVEY.md 8d)",
            "l2_policy": "every field array is 134 MB > 126 MB L2; stage timings additionally flush L2"}


def step_jobs(args, world):
    population = load_population()
    count = len(population) if args.strong else world * args.individuals_per_gpu
    return [population[i % len(population)] for i in range(count)]


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
    jobs = step_jobs(args, world)
    peak, peak_src = measured_peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_stats(values):
        """-> (max over ranks, sum over ranks) of a list of floats."""
        t = torch.tensor(values, dtype=torch.float64, device="cuda")
        if world == 1:
            return t.tolist(), t.tolist()
        mx, sm = t.clone(), t.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        return mx.tolist(), sm.tolist()

    def ordered_jobs(grammar, problem, lo, hi, strings):
        # longest-first by the static cost of one cycle (the number of cycles is not known beforehand)
        cost = [lowering.estimate_cycle_bytes(lowering.lower(grammar.compile(s)[0], lo, hi).ops, problem.dimension)
                for s in strings]
        return [strings[i] for i in sorted(range(len(strings)), key=lambda i: (-cost[i], i))]

    def timed_region(gens, grammar, ordered, lo, hi, steps, problem):
        """`steps` passes over `ordered`, between barriers; device time = max over the rank's streams."""
        wss = [g._workspace for g in gens]
        for g in gens:
            g.h2d_bytes = g.d2h_bytes = 0
        launches0 = sum(w.launch_count() for w in wss)
        accounting = {"bytes": 0.0, "cycle_bytes": lambda program: cycle_bytes_of(program, problem, args.engine_min_level)}
        barrier()
        for w in wss:
            w.mark(0)
        t0 = time.perf_counter()
        cycles, results, counts = 0, None, {}
        for _ in range(steps):
            c, results = evaluate_shard(gens, grammar, ordered, lo, hi, counts, schedule=args.schedule, accounting=accounting)
            cycles += c
        for w in wss:
            w.mark(1)
        for w in wss:
            w.synchronize()
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        dev_ms = max(w.elapsed() for w in wss)          # every stream is marked at the common start and end
        launches = sum(w.launch_count() for w in wss) - launches0
        mx, sm = reduce_stats([dev_ms, wall_ms, float(cycles), float(launches), float(sum(g.h2d_bytes for g in gens)),
                               float(sum(g.d2h_bytes for g in gens)), accounting["bytes"]])
        return {"dev_ms": mx[0], "wall_ms": mx[1], "cycles": sm[2], "launches": sm[3], "h2d": sm[4], "d2h": sm[5],
                "bytes": sm[6], "results": results, "counts": counts}

    def timed_phase(resident):
        made = [make_generator(args, local, resident) for _ in range(max(1, args.streams))]
        gens = [m[0] for m in made]
        gen, problem, lo, hi = made[0]
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
        ordered = ordered_jobs(grammar, problem, lo, hi, jobs)
        us, fs = synthetic_inputs(problem, hi)
        for g in gens:
            g.set_inputs(rhs=fs, u0=us)
        for _ in range(args.warmup):
            evaluate_shard(gens, grammar, ordered, lo, hi, schedule=args.schedule)
        out = timed_region(gens, grammar, ordered, lo, hi, args.steps, problem)
        out.update(gens=gens, problem=problem, lo=lo, hi=hi, grammar=grammar)
        return out

    # clocks are sampled for rank 0's GPU only: one nvidia-smi process per rank every 0.2 s takes the
    # driver's global locks often enough on an 8-GPU box to disturb the other ranks' launches
    with (ClockSampler(local) if rank == 0 else contextlib.nullcontext()) as clocks:
        res = timed_phase(resident=True)
        gens, problem, lo, hi, grammar = res["gens"], res["problem"], res["lo"], res["hi"], res["grammar"]
        stages = stage_rooflines(gens[0], problem, hi, peak) if rank == 0 else None
        strong = None
        if not args.no_extras and not args.strong:
            # configs[1] is a fixed population of 256: one pass over all of it at this N (kernels are warm)
            everything = ordered_jobs(grammar, problem, lo, hi, load_population())
            s = timed_region(gens, grammar, everything, lo, hi, 1, problem)
            dof2 = float((2 ** hi - 1) ** 2)
            strong = {"scaling": "strong", "individuals": len(everything), "ms": s["dev_ms"],
                      "value": dof2 * s["cycles"] / (s["dev_ms"] * 1e-3), "unit": UNIT,
                      "evals_per_s": len(everything) / (s["dev_ms"] * 1e-3), "cycles": s["cycles"],
                      "step_roofline_frac": s["bytes"] / (s["dev_ms"] * 1e-3) / 1e9 / (peak * world)}
        for g in gens:
            g.close()
        e2e_res = timed_phase(resident=False)
        for g in e2e_res["gens"]:
            g.close()
    dof = float((2 ** hi - 1) ** 2)
    dev_ms, cycles = res["dev_ms"], res["cycles"]
    value = dof * cycles / (dev_ms * 1e-3)
    e2e_ms = max(e2e_res["dev_ms"], e2e_res["wall_ms"])
    e2e = dof * e2e_res["cycles"] / (e2e_ms * 1e-3)
    extras = {}
    if not args.no_extras:
        extras["dd3d"] = dd3d_leg(args, rank, world, local, peak)
        if rank == 0:
            extras["small_configs"] = small_config_records(local)
            extras["variable_coefficient"] = variable_coefficient_record(local, peak, args.max_level)
    if rank != 0:
        finish(dist, rank, world)
        return
    results, counts = res["results"], res["counts"]
    converged = sum(1 for r in results if r[2] < 1e99)
    # share of rank 0's timed region spent in each finest-level stage kernel: launches counted from
    # the op lists x the stage's measured launch time; the largest is the dominant kernel
    share = {name: {"launches": k, "est_ms": k * stages[name]["ms"] if name in stages else None,
                    "share": k * stages[name]["ms"] / dev_ms if name in stages else None}
             for name, k in sorted(counts.items())}
    dominant = max((n for n in share if share[n]["share"] is not None), key=lambda n: share[n]["share"], default="rbgs")
    cpu = cpu_baseline(args, world)
    step_gbs = res["bytes"] / (dev_ms * 1e-3) / 1e9
    config = workload_config(args, world)
    config.update({"schedule": args.schedule, "streams_per_gpu": max(1, args.streams),
                   "kernels": "generic" if args.generic_kernels else "specialised+generic"})
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
        "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config,
        "evals_per_s": len(jobs) * args.steps / (dev_ms * 1e-3),
        "cycles_per_step": cycles / args.steps, "converged_in_last_step_rank0": converged,
        "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": e2e_res["h2d"] / args.steps,
                "d2h_bytes_per_step": e2e_res["d2h"] / args.steps, "ms_per_step": e2e_ms / args.steps,
                "input_copy": "synchronous before each solve" if args.no_prefetch else
                              "pinned host -> device staging on a copy stream, one evaluation ahead of the solve"},
        "gpu_launches": int(res["launches"]),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": stages[dominant]["achieved_gbs"], "peak": peak,
                     "unit": "GB/s", "frac": stages[dominant]["frac"], "traffic": captured_traffic(dominant), "peak_source": peak_src,
                     "traffic_source": "profiles/r1_traffic.json (ncu --set full, same kernel and size)",
                     "algorithmic_bytes_per_launch": stages[dominant]["bytes"],
                     "ms_per_launch": stages[dominant]["ms"],
                     # the whole timed region against the same peak: what the step as a whole achieves
                     "step": {"algorithmic_bytes_per_step": res["bytes"] / args.steps, "achieved": step_gbs,
                              "peak": peak * world, "unit": "GB/s", "frac": step_gbs / (peak * world),
                              "bytes": "sum over evaluations of cycles x (op-list bytes of one cycle, SURVEY.md 8d table, "
                                       "+ 16 B/DoF residual norm) + 16 B/DoF initial norm; all ranks"}},
        "stages": stages, "finest_level_share_rank0": share,
        "cpu_baseline": cpu,
        "clocks": clocks.summary(),
    }
    if strong is not None:
        line["strong_scaling"] = strong
    line.update(extras)
    print(json.dumps(line))
    finish(dist, rank, world)


def finish(dist, rank, world):
    """End of a multi-rank run.  Rank 0 still has the CPU legs to time when the others are done: they wait
    on the rendezvous store (a blocking socket read) -- a NCCL barrier would busy-wait a host core each
    and slow the OpenMP team of the CPU leg down (measured: 24 threads 690 ms per cycle against 85)."""
    if world <= 1:
        return
    import datetime
    from torch.distributed import distributed_c10d
    store = distributed_c10d._get_default_store()
    if rank == 0:
        store.set("bench/finished", "1")
    else:
        store.wait(["bench/finished"], datetime.timedelta(seconds=3600))
    dist.destroy_process_group()


def run_dd_only(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    peak, _ = measured_peaks()
    record = dd3d_leg(args, rank, world, local, peak)
    if rank == 0:
        print(json.dumps(record))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def dd3d_leg(args, rank, world, local, peak):
    """BASELINE.json configs[2]: one V(2,1) RB-GS cycle of 3D Poisson 513^3 (the `.exa3` solver block,
    omega = 1.25, CG on level 2), the grid cut into z-slabs over the N GPUs with halo exchange
    (evostencils_b200/domain.py); N = 1: the single-domain engine, one captured CUDA graph per cycle.
    Every rank takes part; -> the record (identical on all ranks)."""
    import torch
    import torch.distributed as dist
    import bench_dd
    from evostencils_b200 import domain, lowering, problems
    problem, level, lo = problems.POISSON_3D, args.dd_level, 2
    ops = bench_dd.cycle_ops(problem, level, lo, 1)
    rhs, u0 = bench_dd.slab_inputs(problem, level)
    dof = float((2 ** level - 1) ** 3)
    bytes_per_cycle = lowering.estimate_cycle_bytes(ops, 3) + 16.0 * dof
    n = 2 ** level + 1
    record = {"workload": "BASELINE.json configs[2]: 3D Poisson 7-point %d^3 fp64, V(2,1) red-black GS omega=1.25, "
                          "CG on level %d, z-slabs over %d GPU(s)" % (n - 1, lo, world),
              "cycles": args.dd_cycles, "algorithmic_bytes_per_cycle": bytes_per_cycle, "scaling": "strong"}
    try:
        if world == 1:
            from evostencils_b200.engine import Workspace
            ws = Workspace(3, lo, level, 1, "f64", local)
            for lv in range(lo, level + 1):
                ws.set_operator(lv, lowering.operator_table(problem.system_operator(lv, 0)))
            ws.set_coarse_solver(problem.solver.cgs, problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
            ws.upload("rhs", level, 0, rhs[0](0, n))
            ws.upload("sol", level, 0, u0[0](0, n))
            r0 = ws.residual_norm()
            cyc = ws.build_cycle(ops)
            ws.run_cycles(cyc, 2)
            ms = ws.run_cycles(cyc, args.dd_cycles) / args.dd_cycles
            r1 = ws.residual_norm()
            ws.close()
            record.update({"driver": "single domain, one CUDA graph per cycle", "halo_bytes_per_cycle": 0.0})
        else:
            comm = domain.SlabComm.from_environment()
            solver = domain.make_solver(problem, lo, level, comm, device=local)
            solver.set_fields(rhs, u0)
            r0 = solver.residual_norm()
            ms, info = solver.time_cycles(ops, args.dd_cycles, warmup=2)
            r1 = solver.residual_norm()
            solver.close()
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = t.item()
            record.update(info)
        gbs = bytes_per_cycle / (ms * 1e-3) / 1e9
        record.update({"ms_per_cycle": ms, "value": dof / (ms * 1e-3), "unit": UNIT,
                       "residual_reduction": r1 / r0 if r0 else None,
                       "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak * world, "unit": "GB/s",
                                    "frac": gbs / (peak * world)}})
    except Exception as exc:                       # the headline line must not die with an auxiliary leg
        record["error"] = "%s: %s" % (type(exc).__name__, exc)
    return record


_CPORT = []


def host_port():
    """oracle/cport, rebuilt once per process for this host's CPU (-march=native; the in-tree build is
    x86-64-v3 so that it runs on whatever box the snapshot lands on)."""
    from oracle import cport
    if not _CPORT:
        cport.load(march="native", force=True)
        _CPORT.append(cport)
    return cport


def _timed(fn):
    t0 = time.perf_counter()
    out = fn()
    return out, (time.perf_counter() - t0) * 1e3


def small_config_records(device):
    """configs[0], [3], [4]: KB..MB grids, latency-bound -- ms per solve and evaluations per second of
    one GPU (replicas ride on the population sharding), the CPU figure of the same solve beside it:
    the C/OpenMP port where it covers the problem (scalar real), else the numpy oracle, bounded to a
    few iterations and scaled to the solve's iteration count (said in `cpu.sample`)."""
    from evostencils_b200 import lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar, v_cycle_string
    from oracle import driver
    cport = host_port()
    out = {}

    def gpu_solve(problem, lo, hi, string, params, samples):
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=4, parameters=params or None)
        program = lowering.lower(grammar.compile(string)[0], lo, hi)
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=device)
        gen.evaluate_program(program, params or {}, evaluation_samples=1)           # warm-up: capture, tables
        t, rho, n = gen.evaluate_program(program, params or {}, evaluation_samples=samples)
        launches = gen.last_launch_count
        gen.close()
        return program, {"ms_per_solve": t, "iterations": n, "convergence_factor": rho, "evals_per_s": 1e3 / t if t > 0 else None}

    try:    # configs[0]: scripts/evaluate_reference_solver.py -- 2D Poisson 5..9, V(2,1) RB-GS w = 1.15, CG on 33^2
        problem, lo, hi = problems.POISSON_2D, 5, 9
        program, rec = gpu_solve(problem, lo, hi, v_cycle_string(hi - lo, 2, 1, 21, "red_black"), None, 20)
        ws = cport.workspace(2, lo, hi)
        for level in range(lo, hi + 1):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.set_coarse_solver("CG", problem.solver.cgs_max_iterations, problem.solver.cgs_target_reduction)
        us, fs = problem.initial_fields(hi)
        best = None
        for _ in range(3):
            ws.upload("sol", hi, 0, us[0])
            ws.upload("rhs", hi, 0, fs[0])
            ws.clear_programs()
            cyc = ws.build_cycle(program.ops, fused=False)
            hist, conv, ms = ws.solve(cyc, problem.solver.target_reduction, problem.solver.max_iterations)
            best = ms if best is None else min(best, ms)
        ws.close()
        rec["cpu"] = {"ms_per_solve": best, "iterations": len(hist) - 1, "kind": "port", "cores": cport.set_threads(len(os.sched_getaffinity(0))),
                      "sample": "whole solve, best of 3, C/OpenMP oracle port"}
        rec["workload"] = "configs[0]: 2D Poisson 513^2, levels 5..9, reference V(2,1) RB-GS w=1.15 (Poisson/2D...exa3:2-15)"
        out["poisson2d_reference_solver"] = rec
    except Exception as exc:
        out["poisson2d_reference_solver"] = {"error": "%s: %s" % (type(exc).__name__, exc)}

    def numpy_leg(problem, program, params, its, n_gpu):
        (h, hist, rho, n), ms = _timed(lambda: driver.solve(problem, program, params=params, max_iterations=its))
        done = max(1, len(hist) - 1)
        scaled = ms / done * n_gpu if n_gpu < 1e99 else None
        return {"ms_per_solve": scaled, "kind": "port", "cores": 1,
                "sample": "numpy oracle, %d iterations timed (%.1f ms each), scaled to the solve's %s iterations"
                          % (done, ms / done, int(n_gpu) if n_gpu < 1e99 else "n/a")}

    try:    # configs[3]: Helmholtz 3..7, k = 80, V(2,1) RB-GS w = 0.6 preconditioner inside BiCGStab
        problem, lo, hi = problems.HELMHOLTZ_2D, 3, 7
        program, rec = gpu_solve(problem, lo, hi, v_cycle_string(hi - lo, 2, 1, 10, "red_black"), {"k": 80.0}, 3)
        rec["cpu"] = numpy_leg(problem, program, {"k": 80.0}, 8, rec["iterations"] + 1)
        rec["workload"] = "configs[3]: 2D Helmholtz 129^2 complex128, levels 3..7, k=80, shifted-Laplacian V(2,1) cycle inside BiCGStab"
        out["helmholtz"] = rec
    except Exception as exc:
        out["helmholtz"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    try:    # configs[4]: elasticity 4..8, `.exa3` baseline V(2,1) collective RB-GS w = 1.25
        problem, lo, hi = problems.LINEAR_ELASTICITY_2D, 4, 8
        program, rec = gpu_solve(problem, lo, hi, v_cycle_string(hi - lo, 2, 1, 23, "red_black"), None, 3)
        rec["cpu"] = numpy_leg(problem, program, None, 2, rec["iterations"])
        rec["workload"] = "configs[4]: 2D linear elasticity 257^2 x 2 fields, levels 4..8, V(2,1) collective RB-GS w=1.25"
        out["elasticity"] = rec
    except Exception as exc:
        out["elasticity"] = {"error": "%s: %s" % (type(exc).__name__, exc)}
    return out


def variable_coefficient_record(device, peak, level=12):
    """north_star's "variable stencils": -div(c grad u), c = exp(10 (x - x^2)(y - y^2)) (stencils/gallery.py:87-131)
    at the size of configs[1].  Stage kernels timed like `stage_rooflines` (24 B/DoF: the coefficients are
    recomputed per node, nothing but u and f is read; a red-black sweep is two in-place colour passes = 48
    B/DoF moved for 24 algorithmic), and one V(2,1) solve through the generator."""
    from evostencils_b200 import ir, lowering, problems
    from evostencils_b200.generator import B200ProgramGenerator
    from evostencils_b200.grammar import Grammar, v_cycle_string
    try:
        problem, lo, hi = problems.POISSON_2D_VARIABLE, 2, level
        gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=hi, device=device)
        gen.initialize_code_generation(lo, hi)
        expression, _ = Grammar(problem, lo, hi).compile(v_cycle_string(hi - lo, 2, 1, 21, "red_black"))
        storages = gen.generate_storage(lo, hi, gen.finest_grid)
        us, fs = synthetic_inputs(problem, hi)          # as the main leg: u0 = 0, f ~ N(0, 1)
        gen.set_inputs(rhs=fs, u0=us)
        gen.generate_and_evaluate(expression, storages, lo, hi, "", evaluation_samples=1)       # warm-up
        t, rho, n = gen.generate_and_evaluate(expression, storages, lo, hi, "", evaluation_samples=1)
        ws = gen._workspace
        op = problem.system_operator(hi, 0)
        dof = float((2 ** hi - 1) ** 2)
        stages = {
            "jacobi": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, hi, 0.8, ir.Single),
            "rbgs": lowering.local_solve_statements(ir.ElementwiseDiagonal(op), op, hi, 1.15, ir.RedBlack),
            "residual": [lowering.Op("residual", hi, dst="res")],
        }
        rec = {}
        for name, ops in stages.items():
            ws.time_ops(ops, 3)
            ws.flush_l2()
            ms = ws.time_ops(ops, 20) / 20
            gbs = 24.0 * dof / (ms * 1e-3) / 1e9
            rec[name] = {"ms": ms, "achieved_gbs": gbs, "frac": gbs / peak, "bytes": 24.0 * dof}
        gen.close()
        return {"workload": "2D variable-coefficient Poisson %d^2 fp64, kappa = 10, levels %d..%d" % (2 ** hi, lo, hi),
                "stages": rec, "solve": {"cycle": "V(2,1) RB-GS w=1.15", "ms": t, "iterations": n, "convergence_factor": rho,
                                         "value": dof * n / (t * 1e-3) if n < 1e99 else None, "unit": UNIT}}
    except Exception as exc:
        return {"error": "%s: %s" % (type(exc).__name__, exc)}


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
    """C/OpenMP oracle port on the host cores: one loop nest per unfused statement -- the granularity
    of the code ExaStencils generates.  A sample = one cycle (+ residual norm) of each of the step's
    first `--cpu-sample-individuals` individuals from the bench's inputs, i.e. the same op mix the GPU
    arm averages over (the GPU arm weights an individual by its cycles to convergence, the sample by
    one cycle each).  The OpenMP team size is set through the library, whatever OMP_NUM_THREADS says
    (torchrun exports OMP_NUM_THREADS=1 to its workers), and reported as OpenMP measured it."""

    def __init__(self, args, world=1):
        from evostencils_b200 import lowering, problems
        from evostencils_b200.grammar import Grammar
        self.cport = cport = host_port()
        self.host_cores = len(os.sched_getaffinity(0))
        problem = problems.POISSON_2D
        self.hi = hi = args.max_level
        lo = hi - args.grammar_depth
        elo = args.engine_min_level
        grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
        strings = step_jobs(args, world)[: args.cpu_sample_individuals]
        self.ws = ws = cport.workspace(2, elo, hi)
        for level in range(elo, hi + 1):
            ws.set_operator(level, lowering.operator_table(problem.system_operator(level, 0)))
        ws.set_coarse_solver("CG", 1000, 1e-12)
        self.us, fs = synthetic_inputs(problem, hi, pinned=False)
        ws.upload("rhs", hi, 0, fs[0])
        default = lambda level: lowering.default_cycle_ops(problem, level, elo)
        self.cycles = [ws.build_cycle(lowering.lower(grammar.compile(s)[0], lo, hi).ops, default, fused=False)
                       for s in strings]
        self.threads = cport.set_threads(self.host_cores)
        ws.upload("sol", hi, 0, self.us[0])
        ws.run_cycles(self.cycles[0], 1)                      # warm-up (page faults)

    def run(self, threads=None):
        used = self.cport.set_threads(threads or self.host_cores)
        ms = 0.0
        for cycle in self.cycles:
            self.ws.upload("sol", self.hi, 0, self.us[0])    # every individual starts from the zero guess (untimed)
            ms += self.ws.run_cycles(cycle, 1)
        n = len(self.cycles)
        dof = float((2 ** self.hi - 1) ** 2)
        return {"value": dof * n / (ms * 1e-3), "unit": UNIT, "cores": used, "kind": "port",
                "sample": "one cycle (+ residual norm) of each of the step's first %d individuals at %d^2, "
                          "C/OpenMP oracle port, unfused statements, %d OpenMP threads (omp_get_num_threads)"
                          % (n, 2 ** self.hi, used),
                "ms_per_cycle": ms / n, "host_cores": self.host_cores}

    def close(self):
        self.ws.close()


def cpu_baseline(args, world=1):
    """All host threads + the reference's own setting of 4 (lib/parallelization_pureOmp.knowledge:3)."""
    port = CpuPort(args, world)
    try:
        out = port.run()
        four = port.run(threads=4)
        out["threads_4"] = {"value": four["value"], "unit": UNIT, "cores": four["cores"], "ms_per_cycle": four["ms_per_cycle"]}
        return out
    finally:
        port.close()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    port = CpuPort(args, world)
    for _ in range(min(args.warmup, 2)):
        port.run()
    t0 = time.perf_counter()
    vals = [port.run() for _ in range(args.steps)]
    wall = (time.perf_counter() - t0) * 1e3
    four = port.run(threads=4)
    port.close()
    value = float(np.mean([v["value"] for v in vals]))
    cpu = dict(vals[-1])
    cpu["value"] = value
    cpu["threads_4"] = {"value": four["value"], "unit": UNIT, "cores": four["cores"], "ms_per_cycle": four["ms_per_cycle"]}
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall / args.steps, "higher_is_better": True, "scaling": "strong" if args.strong else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, world),
            "note": "the reference's own ExaStencils-generated C++ cannot be built here (no java); this is the C/OpenMP "
                    "oracle port (oracle/mg_oracle_c.c); each step is a bounded sample of the workload (cpu_baseline.sample)",
            "cpu_baseline": cpu,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.stages_only:
        run_stages(a)
    elif a.dd_only:
        run_dd_only(a)
    else:
        run_b200(a)

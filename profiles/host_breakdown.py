#!/usr/bin/env python
"""Where one fitness evaluation spends its wall time (config 2): host phases around the solve.
Wraps the phases of B200ProgramGenerator.generate_and_evaluate with perf_counter (device work is
synchronised at each boundary, so this perturbs the overlap; it is a diagnostic, not a bench)."""
import collections
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--individuals", type=int, default=8)
    ap.add_argument("--resident", type=int, default=1)
    a = ap.parse_args()
    from evostencils_b200 import lowering
    from evostencils_b200.grammar import Grammar
    args = bench.parse_args.__wrapped__() if hasattr(bench.parse_args, "__wrapped__") else None
    sys.argv = sys.argv[:1]
    args = bench.parse_args()
    gen, problem, lo, hi = bench.make_generator(args, 0, bool(a.resident))
    grammar = Grammar(problem, lo, hi, maximum_local_system_size=4)
    us, fs = bench.synthetic_inputs(problem, hi)
    gen.set_inputs(rhs=fs, u0=us)
    shard = bench.load_population()[: a.individuals]
    bench.evaluate_shard(gen, grammar, shard, lo, hi)          # warm-up
    t = collections.Counter()
    ws = gen._workspace
    orig = {}

    def wrap(obj, name, key):
        f = getattr(obj, name)
        orig[(obj, name)] = f

        def g(*args, **kw):
            ws.synchronize()
            t0 = time.perf_counter()
            r = f(*args, **kw)
            ws.synchronize()
            t[key] += time.perf_counter() - t0
            return r
        setattr(obj, name, g)

    wrap(ws, "build_cycle", "build_cycle (encode + smoother tables + mg_build_cycle)")
    wrap(ws, "solve", "solve")
    inner_solve = ws.solve

    def solve_dev(*args, **kw):
        r = inner_solve(*args, **kw)
        t["  of which device time between the solve's events (excludes graph capture/instantiate)"] += r[2] * 1e-3
        return r
    ws.solve = solve_dev
    wrap(ws, "clear_programs", "clear_programs")
    wrap(gen, "_load_inputs", "load_inputs")
    wrap(gen, "_configure", "configure")
    lower = lowering.lower

    def timed_lower(*args, **kw):
        t0 = time.perf_counter()
        r = lower(*args, **kw)
        t["lowering.lower"] += time.perf_counter() - t0
        return r
    lowering.lower = timed_lower
    compile_ = grammar.compile

    def timed_compile(s):
        t0 = time.perf_counter()
        r = compile_(s)
        t["grammar.compile"] += time.perf_counter() - t0
        return r
    grammar.compile = timed_compile
    ws.synchronize()
    t0 = time.perf_counter()
    cycles, _ = bench.evaluate_shard(gen, grammar, shard, lo, hi)
    ws.synchronize()
    total = time.perf_counter() - t0
    out = {k: round(v * 1e3 / a.individuals, 3) for k, v in sorted(t.items(), key=lambda kv: -kv[1])}
    out["total ms per individual"] = round(total * 1e3 / a.individuals, 3)
    out["unaccounted"] = round((total - sum(v for k, v in t.items() if not k.startswith("  "))) * 1e3 / a.individuals, 3)
    out["cycles per individual"] = cycles / a.individuals
    print(json.dumps(out, indent=1))
    gen.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Time every specialised stage kernel alone on one level (CUDA events on the engine stream,
L2 flushed before the timed batch) -> one JSON line.  Tunables come from the environment
(MG_*_ROWS, ...), which the engine reads once per process; run one process per setting:

    MG_RBGS_ROWS=45 python profiles/sweep_stages.py --level 12
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--level", type=int, default=12)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    from evostencils_b200 import problems
    from evostencils_b200.generator import B200ProgramGenerator
    problem = problems.POISSON_2D
    lo = max(2, a.level - 3)
    gen = B200ProgramGenerator(problem=problem, min_level=lo, max_level=a.level, device=0)
    gen.initialize_code_generation(lo, a.level)
    ws = gen._ensure_workspace(lo, a.level)
    gen._configure(ws, {})
    us, fs = bench.synthetic_inputs(problem, a.level, pinned=False)
    ws.upload("sol", a.level, 0, us[0])
    ws.upload("rhs", a.level, 0, fs[0])
    peak, _ = bench.measured_peaks()
    stages = bench.stage_rooflines(gen, problem, a.level, peak)
    env = {k: v for k, v in os.environ.items() if k.startswith("MG_")}
    print(json.dumps({"tag": a.tag, "level": a.level, "env": env,
                      "stages": {k: {"ms": round(v["ms"], 5), "frac": round(v["frac"], 4)} for k, v in stages.items()}}))
    gen.close()


if __name__ == "__main__":
    main()

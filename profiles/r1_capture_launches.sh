#!/bin/bash
# Round 1, call A: the default bench line (no profiler), then the ncu launch list of the same
# program on a bounded window (2500 launches = the first individuals' evaluations; ncu replays
# every kernel, a whole default run has ~45k launches).  Run under gpurun from the repo root.
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/r1_bench_default.json 2> gpurun_out/r1_bench_default.err || exit 1
python bench.py --steps 1 --warmup 1 --individuals-per-gpu 2 > gpurun_out/r1_bench_short.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv \
    --log-file gpurun_out/r1_launch_list.csv \
    python bench.py --steps 1 --warmup 1 --individuals-per-gpu 2 > gpurun_out/r1_launch_list.log 2>&1
echo "ncu rc=$?"
tail -c 600 gpurun_out/r1_bench_default.json

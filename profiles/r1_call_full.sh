#!/bin/bash
# gpurun call: whole GPU test-suite, config 3 cycle profile and single-GPU bench_dd line
mkdir -p gpurun_out
( time timeout -k 10 1200 python -m pytest tests -x -q -m gpu ) > gpurun_out/pytest_gpu.log 2>&1
tail -4 gpurun_out/pytest_gpu.log
timeout -k 10 200 python profiles/profile_3d.py --level 9 > gpurun_out/p3d_l9.json 2> gpurun_out/p3d_l9.err; cat gpurun_out/p3d_l9.json
timeout -k 10 300 python bench_dd.py > gpurun_out/bench_dd_1.json 2> gpurun_out/bench_dd_1.err; cat gpurun_out/bench_dd_1.json | cut -c1-700

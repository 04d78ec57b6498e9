#!/bin/bash
# gpurun call (round 2): 3D stage kernels at 513^3 (event timing) + optional ncu captures
mkdir -p gpurun_out/f
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_sizes.py -x -q -m gpu -k "3d or one_pass or 513 or 257" 2>&1 | tail -3
echo "rbgs3d:     $(timeout -k 10 120 python profiles/time_rbgs3d.py --level 9 2>&1 | tail -1 | cut -c1-90)"
echo "restrict3d: $(timeout -k 10 120 python profiles/time_restrict3d.py --level 9 2>&1 | tail -1 | cut -c1-90)"
echo "cycle:      $(timeout -k 10 200 python bench.py --dd-only 2>/dev/null | tail -1 | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["ms_per_cycle"], d["roofline"]["frac"])')"
if [ -n "$NCU" ]; then
timeout -k 10 300 ncu --set full --clock-control none --import-source on -k regex:k_rbgs3d -c 1 -f -o gpurun_out/f/rbgs3d_513 python profiles/time_rbgs3d.py --level 9 --reps 1 > gpurun_out/f/ncu_rbgs3d.log 2>&1; echo "ncu rbgs rc=$?"
timeout -k 10 300 ncu --set full --clock-control none --import-source on -k regex:k_resrestrict3d -c 1 -f -o gpurun_out/f/restrict3d_513 python profiles/time_restrict3d.py --level 9 --reps 1 > gpurun_out/f/ncu_restrict3d.log 2>&1; echo "ncu restrict rc=$?"
fi

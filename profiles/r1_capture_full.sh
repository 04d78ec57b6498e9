#!/bin/bash
# Round 1, call B: one `--set full` capture of the dominant kernel (fused two-colour RB-GS at
# 4096^2) and of the Jacobi/residual tile kernel, two launches each, inside the bench program.
set -u
mkdir -p gpurun_out
python bench.py --steps 1 --warmup 1 --individuals-per-gpu 1 > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_rbgs2d|k_star2d' \
    --launch-skip 40 -c 6 -o gpurun_out/r1_full_top -f \
    python bench.py --steps 1 --warmup 1 --individuals-per-gpu 1 > gpurun_out/r1_full_top.log 2>&1
echo "ncu rc=$?"
ls -la gpurun_out

#!/bin/bash
# Round 1, call B: `--set full` captures of the finest-level (4096^2) stage kernels, launched by
# `bench.py --stages-only` (the same stage_rooflines() the bench line's roofline comes from;
# 1 warm-up + 1 timed launch per stage under ncu, so 10 captured launches).
set -u
mkdir -p gpurun_out
python bench.py --stages-only --stage-reps 1 > gpurun_out/r1_stages_plain1.json 2>&1 || { cat gpurun_out/r1_stages_plain1.json; exit 1; }
python bench.py --stages-only > gpurun_out/r1_stages_plain20.json 2>&1 || exit 1
ncu --set full --clock-control none --import-source on \
    -k regex:'k_rbgs2d|k_star2d|k_resrestrict2d|k_prolong2d' -c 12 -o gpurun_out/r1_full_stages -f \
    python bench.py --stages-only --stage-reps 1 > gpurun_out/r1_full_stages.log 2>&1
echo "ncu rc=$?"
cat gpurun_out/r1_stages_plain20.json
ls -la gpurun_out

#!/bin/bash
# Round 1, final capture (one gpurun call, one GPU):
#  A  default bench line (no profiler)
#  B  ncu launch list of the bench program on a bounded window (2 individuals, 2500 launches)
#  C  ncu --set full of the six finest-level 2D stage kernels launched by `bench.py --stages-only`
#  D  ncu --set full of the two TMA-staged 3D kernels at 257^3, and their event timings at 513^3
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/r1_bench_default.json 2> gpurun_out/r1_bench_default.err || { tail -5 gpurun_out/r1_bench_default.err; exit 1; }
python bench.py --stages-only > gpurun_out/r1_stages_plain20.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r1_launch_list.csv \
    python bench.py --steps 1 --warmup 1 --individuals-per-gpu 2 > gpurun_out/r1_launch_list.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on \
    -k regex:'k_rbgs2d|k_star2d|k_resrestrict2d|k_prolong2d|k_tile2d' -c 14 -o gpurun_out/r1_full_stages -f \
    python bench.py --stages-only --stage-reps 1 > gpurun_out/r1_full_stages.log 2>&1
echo "full 2D rc=$?"
python profiles/time_rbgs3d.py > gpurun_out/r1_rbgs3d_513.json 2>&1
python profiles/time_restrict3d.py > gpurun_out/r1_restrict3d_513.json 2>&1
MG_RBGS3D=0 python profiles/time_rbgs3d.py > gpurun_out/r1_rbgs3d_513_two_pass.json 2>&1
MG_RESTRICT3D=0 python profiles/time_restrict3d.py > gpurun_out/r1_restrict3d_513_unfused.json 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_rbgs3d -c 1 -f -o gpurun_out/r1_full_rbgs3d \
    python profiles/time_rbgs3d.py --level 8 --reps 1 > gpurun_out/r1_full_rbgs3d.log 2>&1
echo "full rbgs3d rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_resrestrict3d -c 1 -f -o gpurun_out/r1_full_restrict3d \
    python profiles/time_restrict3d.py --level 8 --reps 1 > gpurun_out/r1_full_restrict3d.log 2>&1
echo "full restrict3d rc=$?"
python profiles/profile_3d.py --level 9 > gpurun_out/r1_profile_3d_l9.json 2>/dev/null
cat gpurun_out/r1_rbgs3d_513.json gpurun_out/r1_restrict3d_513.json gpurun_out/r1_profile_3d_l9.json
tail -c 400 gpurun_out/r1_bench_default.json

#!/bin/bash
# gpurun call: parity of the one-pass 3D RB-GS kernel, its effect on the 513^3 cycle, then the bench line
mkdir -p gpurun_out
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "3d or 3D or bitwise" > gpurun_out/t3d.log 2>&1
echo "pytest-3d rc=$?" >> gpurun_out/t3d.log
tail -5 gpurun_out/t3d.log
for lvl in 8 9; do
  for on in 1 0; do
    MG_RBGS3D=$on timeout -k 10 200 python profiles/profile_3d.py --level $lvl > gpurun_out/p3d_l${lvl}_fused${on}.json 2> gpurun_out/p3d_l${lvl}_fused${on}.err
    echo "level $lvl one-pass=$on: $(cat gpurun_out/p3d_l${lvl}_fused${on}.json)"
  done
done
timeout -k 10 600 python bench.py > gpurun_out/bench_p32.json 2> gpurun_out/bench_p32.err
echo "bench rc=$?"; cat gpurun_out/bench_p32.json | cut -c1-1500

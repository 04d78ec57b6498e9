#!/bin/bash
# gpurun call: variants of the one-pass 3D RB-GS kernel at 513^3 (event timing), optional ncu capture
mkdir -p gpurun_out
timeout -k 10 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "3d_stages or one_pass" 2>&1 | tail -3
echo "two colour passes: $(MG_RBGS3D=0 timeout -k 10 120 python profiles/time_rbgs3d.py 2>&1 | tail -1 | cut -c1-80)"
for cfg in ${CFGS:-0 1 2 3}; do
  for zc in ${ZCS:-32 64}; do
    for pf in ${PFS:-0}; do
      echo "cfg $cfg zchunk $zc pf $pf: $(MG_RBGS3D_CFG=$cfg MG_RBGS3D_ZCHUNK=$zc MG_RBGS3D_PF=$pf timeout -k 10 120 python profiles/time_rbgs3d.py --level ${LEVEL:-9} 2>&1 | tail -1 | cut -c1-80)"
    done
  done
done
if [ -n "$NCU_CFG" ]; then
MG_RBGS3D_CFG=$NCU_CFG timeout -k 10 300 ncu --set full --clock-control none --import-source on -k regex:k_rbgs3d -c 1 -f -o gpurun_out/rbgs3d_cfg$NCU_CFG python profiles/time_rbgs3d.py --level 8 --reps 1 > gpurun_out/ncu_rbgs3d.log 2>&1
echo "ncu rc=$?"
fi

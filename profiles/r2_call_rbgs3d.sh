#!/bin/bash
# gpurun call (round 2): k_rbgs3d (one row per warp) against k_rbgs3d2 (two rows per warp) at 513^3
mkdir -p gpurun_out/g
timeout -k 10 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_config_sizes.py tests/test_domain.py -x -q -m gpu -k "3d or one_pass or 513 or 257 or slab" 2>&1 | tail -3
echo "v1 default: $(MG_RBGS3D_V=1 timeout -k 10 120 python profiles/time_rbgs3d.py --level 9 2>&1 | tail -1 | cut -c1-90)"
for cfg in ${CFGS:-0 1 2 3 4}; do
  echo "v2 cfg $cfg: $(MG_RBGS3D_CFG=$cfg timeout -k 10 120 python profiles/time_rbgs3d.py --level 9 2>&1 | tail -1 | cut -c1-90)"
done
echo "v2 cfg 0 level 8: $(timeout -k 10 120 python profiles/time_rbgs3d.py --level 8 2>&1 | tail -1 | cut -c1-90)"
echo "v1 level 8: $(MG_RBGS3D_V=1 timeout -k 10 120 python profiles/time_rbgs3d.py --level 8 2>&1 | tail -1 | cut -c1-90)"
if [ -n "$NCU" ]; then
MG_RBGS3D_CFG=$NCU timeout -k 10 300 ncu --set full --clock-control none --import-source on -k regex:k_rbgs3d -c 1 -f -o gpurun_out/g/rbgs3d2_513 python profiles/time_rbgs3d.py --level 9 --reps 1 > gpurun_out/g/ncu_rbgs3d2.log 2>&1; echo "ncu rc=$?"
fi

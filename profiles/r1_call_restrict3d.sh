#!/bin/bash
# gpurun call: parity and timing of the fused 3D residual + restriction kernel
mkdir -p gpurun_out
timeout -k 10 200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "3d_stages or one_pass" 2>&1 | tail -5
echo "residual + generic restriction: $(MG_RESTRICT3D=0 timeout -k 10 120 python profiles/time_restrict3d.py 2>&1 | tail -1 | cut -c1-90)"
for cfg in 0 1 2 3; do
  for zc in 32 64; do
    echo "cfg $cfg zchunk $zc: $(MG_RESTRICT3D_CFG=$cfg MG_RESTRICT3D_ZCHUNK=$zc timeout -k 10 120 python profiles/time_restrict3d.py 2>&1 | tail -1 | cut -c1-90)"
  done
done

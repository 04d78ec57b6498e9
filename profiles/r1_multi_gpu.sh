#!/bin/bash
# Round 1 multi-GPU record (run with `gpurun --gpus 8`): config 3 (513^3, z-slab domain decomposition,
# NCCL halo exchange) at N = 1, 2, 4, 8 and the config-2 population bench at N = 8 and 2.
set -u
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node "$1" --master-addr 127.0.0.1 --master-port "$2" "${@:3}"; }
python bench_dd.py --gpus 1 --level 9 --cycles 5 --warmup 2 2>/dev/null | grep '^{' > gpurun_out/r1_dd_n1.json
for n in 2 4 8; do
  run $n $((29500 + n)) bench_dd.py --gpus $n --level 9 --cycles 5 --warmup 2 2>/dev/null | grep '^{' > gpurun_out/r1_dd_n$n.json
done
run 8 29610 bench.py --gpus 8 --steps 2 --warmup 3 2>/dev/null | grep '^{' > gpurun_out/r1_bench_n8.json
run 2 29611 bench.py --gpus 2 --steps 2 --warmup 3 2>/dev/null | grep '^{' > gpurun_out/r1_bench_n2.json
for f in gpurun_out/r1_dd_n*.json gpurun_out/r1_bench_n*.json; do echo "$f: $(cut -c1-330 $f)"; done

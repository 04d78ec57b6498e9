#!/bin/bash
# gpurun --gpus 2 call: population bench, self-scheduled (default) and strided, 2 GPUs
mkdir -p gpurun_out
for sched in dynamic static; do
  timeout -k 10 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 2 --steps 3 --warmup 3 --schedule $sched > gpurun_out/bench_n2_$sched.json 2> gpurun_out/bench_n2_$sched.err
  echo "$sched rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_n2_$sched.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "evals_per_s", "cycles_per_step", "n_gpus")}, d["e2e"]["value"])
except Exception as e:
    print("no line", e)
PY
done

#!/bin/bash
# gpurun call: population bench with 1, 2, 3 evaluation streams per GPU
mkdir -p gpurun_out
for k in ${STREAMS:-1 2 3}; do
  timeout -k 10 400 python bench.py --steps 2 --warmup 3 --streams $k > gpurun_out/bench_streams$k.json 2> gpurun_out/bench_streams$k.err
  echo "streams $k rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_streams$k.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step", "evals_per_s", "cycles_per_step", "gpu_launches")}, "e2e", d["e2e"]["value"], d["clocks"])
except Exception as e:
    print("no line", e); print(open("gpurun_out/bench_streams$k.err").read()[-1500:])
PY
done

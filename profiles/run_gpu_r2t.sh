#!/bin/bash
mkdir -p gpurun_out
for thr in "" 1 7; do
  export CERES_HOST_FILL_THREADS=$thr; [ -z "$thr" ] && unset CERES_HOST_FILL_THREADS
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2955${thr:-0} bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.1 > gpurun_out/bench_n8_fill_${thr:-auto}.json 2> gpurun_out/bench_n8_fill.err
  python - <<PY
import json
d = json.load(open("gpurun_out/bench_n8_fill_${thr:-auto}.json"))
print("fill threads '${thr:-auto}': e2e", round(d["e2e"]["ms_per_step"], 3), "retained", round(d["e2e_retained"]["ms_per_step"], 3), "ppm8", round(d["e2e_ppm8"]["ms_per_step"], 3), "device", round(d["ms_per_step"], 4))
PY
done

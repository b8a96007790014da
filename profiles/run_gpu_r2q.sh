#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2q.log; cat gpurun_out/pytest_gpu_r2q.log
NG=$(nvidia-smi -L | wc -l)
if [ "$NG" -ge 2 ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.3 > gpurun_out/bench_n2_r2q.json 2> gpurun_out/bench_n2_r2q.err; echo "bench exit $?"; tail -3 gpurun_out/bench_n2_r2q.err
  python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n2_r2q.json"))
print({k: d.get(k) for k in ("n_gpus", "value", "ms_per_step")}, "e2e", d["e2e"]["ms_per_step"], d["e2e"]["d2h_bytes_per_step"], "retained", d["e2e_retained"]["ms_per_step"], "ppm8", d["e2e_ppm8"]["ms_per_step"])
PY
fi

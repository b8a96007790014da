#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2j.log; cat gpurun_out/pytest_gpu_r2j.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_r2j.json 2> gpurun_out/bench_n1_r2j.err; echo "bench exit $?"; tail -3 gpurun_out/bench_n1_r2j.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1_r2j.json"))
for k in ("value", "ms_per_step", "gpu_launches", "kernels_ms"): print(k, d[k])
print("sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "floor", d["e2e"]["floor_ms"], "pageable", d["e2e_pageable"]["ms_per_step"], "ppm8", d["e2e_ppm8"]["ms_per_step"], "ppm16", d["e2e_ppm8"]["ppm16"]["ms_per_step"])
print("shim", json.dumps(d.get("e2e_shim"), indent=1))
print("cpu", d.get("cpu_baseline"))
PY

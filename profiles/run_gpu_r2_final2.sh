#!/bin/bash
# Round-2 confirmation session on the final library: the static-scratch entry point alone in a fresh process, smoke, every
# GPU test (test_zz_no_alloc.py last), both bench arms.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 > gpurun_out/pytest_zz_alone.log; cat gpurun_out/pytest_zz_alone.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final2.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke_final2.log; tail -2 gpurun_out/smoke_final2.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu_final2.log; cat gpurun_out/pytest_gpu_final2.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_n1_final2.json 2> gpurun_out/bench_ref_n1_final2.err; echo "ref arm exit $?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_final2.json 2> gpurun_out/bench_n1_final2.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1_final2.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1_final2.json"))
for k in ("value", "ms_per_step", "gpu_launches"): print(k, d[k])
print("sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "floor", d["e2e"]["floor_ms"], "retained", d["e2e_retained"]["ms_per_step"], "pageable", d["e2e_pageable"]["ms_per_step"], "ppm8", d["e2e_ppm8"]["ms_per_step"])
print("shim", {k: (v["ms_per_step"] if isinstance(v, dict) else v) for k, v in d.get("e2e_shim", {}).items() if k not in ("path", "unit")})
print("roofline", d["roofline"]["frac"], d["roofline"]["issue"], "clocks", d["clocks"])
r = json.load(open("gpurun_out/bench_ref_n1_final2.json")); print("reference arm", r["value"], r["ms_per_step"], r.get("cpu_baseline", {}).get("cores"))
PY

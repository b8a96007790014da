#!/bin/bash
# Session after the builder change (top-down binned SAH is the default GPU build): smoke, every GPU test, the sanitizer substitute
# over the extended driver (static scratch entry, double shards, all three builders), both bench arms.
mkdir -p gpurun_out
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final3.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke_final3.log; tail -2 gpurun_out/smoke_final3.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/pytest_gpu_final3.log; cat gpurun_out/pytest_gpu_final3.log
bash profiles/run_sanitizers.sh gpurun_out 2>&1 | tail -5
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_n1_final3.json 2> gpurun_out/bench_ref_n1_final3.err; echo "ref arm exit $?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_final3.json 2> gpurun_out/bench_n1_final3.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1_final3.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1_final3.json"))
for k in ("value", "ms_per_step", "gpu_launches"): print(k, d[k])
print("sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "floor", d["e2e"]["floor_ms"], "retained", d["e2e_retained"]["ms_per_step"], "pageable", d["e2e_pageable"]["ms_per_step"], "ppm8", d["e2e_ppm8"]["ms_per_step"])
print("device_built_tree", {k: v for k, v in d.get("device_built_tree", {}).items() if k not in ("note", "builder")})
r = json.load(open("gpurun_out/bench_ref_n1_final3.json")); print("reference arm", r["value"], r["ms_per_step"], r.get("cpu_baseline", {}).get("cores"))
PY

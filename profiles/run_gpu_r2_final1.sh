#!/bin/bash
# Round-2 final single-GPU session: smoke, every GPU test, both bench arms, the sanitizer substitute, sequence and grid workloads.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -6 > gpurun_out/pytest_gpu_final.log; cat gpurun_out/pytest_gpu_final.log
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_ref_n1_final.json 2> gpurun_out/bench_ref_n1_final.err; echo "ref arm exit $?"
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1_final.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1_final.json"))
for k in ("value", "ms_per_step", "gpu_launches", "kernels_ms"): print(k, d[k])
print("sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "floor", d["e2e"]["floor_ms"], "retained", d["e2e_retained"]["ms_per_step"], "pageable", d["e2e_pageable"]["ms_per_step"], "ppm8", d["e2e_ppm8"]["ms_per_step"], "ppm16", d["e2e_ppm8"]["ppm16"]["ms_per_step"])
print("shim", {k: (v["ms_per_step"] if isinstance(v, dict) else v) for k, v in d.get("e2e_shim", {}).items() if k not in ("path", "unit")})
print("issue", d["roofline"]["issue"]); print("cpu", d.get("cpu_baseline"))
r = json.load(open("gpurun_out/bench_ref_n1_final.json")); print("reference arm", r["value"], r["ms_per_step"], r.get("cpu_baseline", {}).get("cores"))
PY
for wl in dragon_1080p_primary dragon_1080p_shadow bunny_1080p_primary bunny_1080p_shadow grid448_8k_shadow; do
  timeout 600 python bench.py --workload $wl --steps 10 --warmup 3 --sustain-s 0.3 --verify > gpurun_out/bench_${wl}_final.json 2> gpurun_out/bench_${wl}_final.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_${wl}_final.json"))
    print("$wl", round(d["value"]), "Mrays/s", round(d["ms_per_step"], 3), "ms; e2e", round(d["e2e"]["ms_per_step"], 3), "; cpu", round(d["cpu_baseline"]["value"], 1), "; verified", d.get("verified", {}).get("id_mismatches"), d.get("verified", {}).get("counters_equal"), d.get("verified", {}).get("max_abs_rgb_diff"))
except Exception as e:
    print("$wl failed:", e)
PY
done
timeout 600 python bench.py --workload dragon_turntable_4k --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_seq4k_n1_final.json 2> gpurun_out/bench_seq4k_n1_final.err; cut -c1-400 gpurun_out/bench_seq4k_n1_final.json
bash profiles/run_sanitizers.sh gpurun_out 2>&1 | tail -4

#!/bin/bash
# Both bench arms under torchrun at N = 2 on the final library (the driver's launch line), and the double path's shards.
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_ref_n2_final4.json 2> gpurun_out/bench_ref_n2_final4.err; echo "ref arm N=2 exit $?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29656 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2_final4.json 2> gpurun_out/bench_n2_final4.err; echo "bench N=2 exit $?"; tail -2 gpurun_out/bench_n2_final4.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n2_final4.json"))
print({k: d.get(k) for k in ("n_gpus", "value", "ms_per_step", "gpu_launches")}, "sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "ppm8", d.get("e2e_ppm8", {}).get("ms_per_step"), "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
r = json.load(open("gpurun_out/bench_ref_n2_final4.json")); print("reference arm", r.get("value"), r.get("ms_per_step"))
PY

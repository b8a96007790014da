#!/bin/bash
# Final library on all GPUs of the box: the headline bench under the driver's launch line, then the double path's shards.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29671 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n${N}_final5.json 2> gpurun_out/bench_n${N}_final5.err; echo "bench N=$N exit $?"
python - <<PY
import json
d = json.load(open("gpurun_out/bench_n${N}_final5.json"))
print({k: d.get(k) for k in ("n_gpus", "value", "ms_per_step", "gpu_launches")}, "sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "ppm8", d.get("e2e_ppm8", {}).get("ms_per_step"), "clocks", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29672 profiles/probes/f64_shared.py 2>&1 | grep "^N=" | tee gpurun_out/f64_shared_n$N.log

#!/bin/bash
# Round-2 scaling session on one multi-GPU box: the headline frame at N = 8, 4, 2 (and 1 with FULL=1), then the per-rank timeline at N = 8.
# Run with: gpurun --gpus 8 --timeout 900 -- 'bash profiles/run_gpu_r2_scale.sh'
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
TAG=${TAG:-r2}
run() {  # n port name args...
  local n=$1 port=$2 name=$3; shift 3
  if [ "$n" -gt "$NG" ]; then echo "$name skipped ($NG GPUs)"; return; fi
  if [ "$n" -eq 1 ]; then
    timeout 300 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@" \
      > gpurun_out/$name.json 2> gpurun_out/$name.err
  fi
  echo "$name exit $?"; python - <<PY
import json
try:
    d = json.load(open("gpurun_out/$name.json"))
    print({k: d.get(k) for k in ("n_gpus", "value", "ms_per_step", "kernels_ms_per_rank")}, "sustained", d["sustained"]["ms_per_step"], "e2e", d["e2e"]["ms_per_step"])
except Exception as e:
    print("no line:", e); print(open("gpurun_out/$name.err").read()[-1500:])
PY
}
run 8 29521 bench_n8_$TAG --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.5
run 4 29522 bench_n4_$TAG --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.5
run 2 29523 bench_n2_$TAG --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.5
[ -n "$FULL" ] && run 1 29524 bench_n1_$TAG --steps 20 --warmup 3 --no-cpu-baseline --sustain-s 0.5
if [ "$NG" -ge 8 ]; then
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 profiles/probes/peer_timeline.py 2>&1 | grep -v Warning | grep "rank\|N 8" > gpurun_out/peer_timeline_$TAG.log
  cat gpurun_out/peer_timeline_$TAG.log
fi
if [ -n "$EXTRA" ] && [ "$NG" -ge 8 ]; then
  run 8 29525 bench_seq4k_n8_$TAG --workload dragon_turntable_4k --steps 3 --warmup 3 --no-cpu-baseline
  run 8 29526 bench_grid448_8k_n8_$TAG --workload grid448_8k_shadow --steps 5 --warmup 3 --no-cpu-baseline --sustain-s 0.3
fi

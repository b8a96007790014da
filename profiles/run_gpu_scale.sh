#!/bin/bash
# Scaling session on one 8-GPU box: the headline frame at N = 4 and 8, the 4K turntable and the 10.5 M-triangle 8K scene at N = 8.
# Run with: gpurun --gpus 8 --timeout 1200 -- 'bash profiles/run_gpu_scale.sh'
mkdir -p gpurun_out
run() {  # n port name args...
  local n=$1 port=$2 name=$3; shift 3
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n "$@" \
      > gpurun_out/$name.json 2> gpurun_out/$name.err
  echo "$name exit $?"; cut -c1-420 gpurun_out/$name.json
}
run 8 29521 bench_n8_final --steps 20 --warmup 3 --no-cpu-baseline
run 4 29522 bench_n4_final --steps 20 --warmup 3 --no-cpu-baseline
run 8 29523 bench_seq4k_n8_final --workload dragon_turntable_4k --steps 3 --warmup 3 --no-cpu-baseline
run 8 29524 bench_grid448_8k_n8_final --workload grid448_8k_shadow --steps 5 --warmup 3 --no-cpu-baseline

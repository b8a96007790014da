#!/bin/bash
# Round-2 session B: bench.py (new line format) at N = 1 and, when the box has 2 GPUs, N = 2.
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?" >> gpurun_out/bench_n1.err
tail -3 gpurun_out/bench_n1.err; cat gpurun_out/bench_n1.json
if [ "$N" -ge 2 ]; then
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench exit $?" >> gpurun_out/bench_n2.err
  tail -3 gpurun_out/bench_n2.err; cat gpurun_out/bench_n2.json
fi

#!/bin/bash
# Double path with row-block shards: its tests, then the shared-frame probe on 1 and 2 GPUs.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_f64.py -m gpu -q 2>&1 | tail -8
timeout 300 python profiles/probes/f64_shared.py 2>&1 | tail -4 | tee gpurun_out/f64_shared_n1.log
N=$(nvidia-smi -L | wc -l)
if [ "$N" -ge 2 ]; then
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 profiles/probes/f64_shared.py 2>&1 | grep -v "^W\|^\*" | tail -6 | tee gpurun_out/f64_shared_n$N.log
fi

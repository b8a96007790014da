#!/bin/bash
# A/B: newest traversal-stack entry kept in a register (the pop feeds the next fetch without a shared-memory load).
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" topreg; do
  echo "== variant '${v:-product}' rep $rep"
  CERES_B200_VARIANT=$v timeout 300 python profiles/kbench.py --workload dragon_4k_shadow --frames 30 2>&1 | tail -1
  CERES_B200_VARIANT=$v timeout 300 python profiles/probes/shard_alone.py --shards 8 --frames 30 2>&1 | tail -2
done
done 2>&1 | tee gpurun_out/topreg_ab.log
CERES_B200_VARIANT=topreg timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2

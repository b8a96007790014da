#!/bin/bash
# A/B of the cooperative triangle phase (every lane of the warp takes one (ray, triangle) job of the step): whole 4K frame (kbench
# checks the hit-id hash and the counters) and one rank's share of 8, each library in its own process, two interleaved repeats.
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" coop coop7; do
  echo "== variant '${v:-product}' rep $rep"
  CERES_B200_VARIANT=$v timeout 300 python profiles/kbench.py --workload dragon_4k_shadow --frames 30 2>&1 | tail -1
  CERES_B200_VARIANT=$v timeout 300 python profiles/probes/shard_alone.py --shards 8 --frames 30 2>&1 | tail -2
done
done 2>&1 | tee gpurun_out/coop_ab.log
CERES_B200_VARIANT=coop timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3

#!/bin/bash
# A/B of the prefetch experiments (children's pairs at the top of a step; the next triangle of a leaf): whole 4K frame (kbench:
# hit-id hash checked) and one rank's share of 8 alone on the GPU, each library in its own process, twice interleaved.
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" pfc pft pfb; do
  echo "== variant '${v:-product}' rep $rep"
  CERES_B200_VARIANT=$v timeout 300 python profiles/kbench.py --workload dragon_4k_shadow --frames 30 2>&1 | tail -2
  CERES_B200_VARIANT=$v timeout 300 python profiles/probes/shard_alone.py --shards 8 --frames 30 2>&1 | tail -2
done
done 2>&1 | tee gpurun_out/prefetch_ab.log

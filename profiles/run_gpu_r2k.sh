#!/bin/bash
mkdir -p gpurun_out
for v in "" prefetch stcs both; do
  echo "== variant '$v'"
  CERES_B200_VARIANT=$v timeout 300 python profiles/kbench.py --share 1 --frames 20 2>&1 | tail -1
  CERES_B200_VARIANT=$v timeout 300 python profiles/probes/shard_alone.py --shards 8 --share 2 2>&1 | grep "rank  0"
done

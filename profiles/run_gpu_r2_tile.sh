#!/bin/bash
# Primary rays per warp and cursor grab (32 = an 8x4 tile, 16, 8): whole 4K frame and one rank's share of 8 / 4 alone.
mkdir -p gpurun_out
for t in 32 16 8; do
  echo "== tile lanes $t"
  timeout 300 python profiles/kbench.py --workload dragon_4k_shadow --frames 30 --block $t 2>&1 | tail -1
  timeout 300 python profiles/probes/shard_alone.py --shards 8,4 --frames 30 --tile $t 2>&1 | tail -4
done 2>&1 | tee gpurun_out/tile_lanes_ab.log

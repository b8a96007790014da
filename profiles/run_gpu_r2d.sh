#!/bin/bash
mkdir -p gpurun_out
timeout 300 python profiles/probes/timeline.py --shards 8 --share 1,2 --rank 0 > gpurun_out/timeline_n8.log 2>&1; cat gpurun_out/timeline_n8.log
timeout 300 python profiles/probes/timeline.py --shards 8 --share 2 --rank 7 >> gpurun_out/timeline_n8.log 2>&1; tail -22 gpurun_out/timeline_n8.log
timeout 300 python profiles/probes/shard_alone.py --shards 8 --all-ranks --block-rows 16,4 > gpurun_out/shard_alone_all.log 2>&1; cat gpurun_out/shard_alone_all.log

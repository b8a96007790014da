#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2f.log; cat gpurun_out/pytest_gpu_r2f.log
timeout 300 python profiles/kbench.py --share 1,2 --block 0,16,8 --frames 20 > gpurun_out/kbench_r2f.log 2>&1; cat gpurun_out/kbench_r2f.log
timeout 300 python profiles/probes/shard_alone.py --shards 8,4,2 --chunk 32,16,8 --share 2 2>&1 | grep "rank  0" > gpurun_out/shard_alone_r2f.log; cat gpurun_out/shard_alone_r2f.log
timeout 300 python profiles/probes/shard_alone.py --shards 8 --chunk 32,16 --share 1 2>&1 | grep "rank  0" >> gpurun_out/shard_alone_r2f.log; tail -2 gpurun_out/shard_alone_r2f.log

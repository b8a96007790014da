#!/bin/bash
mkdir -p gpurun_out
timeout 300 python profiles/kbench.py --share 1,2 --records 32,16 --frames 20 > gpurun_out/kbench_r2g.log 2>&1; cat gpurun_out/kbench_r2g.log
timeout 300 python profiles/probes/shard_alone.py --shards 8,4,2 --chunk 32,16 --share 2 2>&1 | grep "rank  0" > gpurun_out/shard_alone_r2g.log; cat gpurun_out/shard_alone_r2g.log
timeout 300 python profiles/probes/shard_alone.py --shards 8,4,2 --share 1 2>&1 | grep "rank  0" >> gpurun_out/shard_alone_r2g.log; tail -3 gpurun_out/shard_alone_r2g.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_peer.py -m gpu -q -x 2>&1 | tail -3

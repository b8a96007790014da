#!/bin/bash
# Round-2 session C: shadow-subtree sharing -- parity tests, kernel A/B (off / on / period and threshold sweep), per-rank shares, timeline.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "sharing or shared or statistics or shadow_order" 2>&1 | tail -8 > gpurun_out/pytest_share.log; cat gpurun_out/pytest_share.log
timeout 300 python profiles/kbench.py --share 1,2,2:1:1,2:2:4,2:2:8,2:4:4,2:4:12,2:8:6,2:3:6 --frames 20 > gpurun_out/kbench_share.log 2>&1; cat gpurun_out/kbench_share.log
timeout 300 python profiles/probes/timeline.py > gpurun_out/timeline.log 2>&1; cat gpurun_out/timeline.log
timeout 300 python profiles/probes/shard_alone.py > gpurun_out/shard_alone.log 2>&1; cat gpurun_out/shard_alone.log

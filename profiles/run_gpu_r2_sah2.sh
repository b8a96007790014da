#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/sah_trace_agg.log 2>&1; tail -29 gpurun_out/sah_trace_agg.log
timeout 600 python profiles/lbvh_probe.py dragon_4k_shadow 0:8:8:1.5,2:0:8:1.5,2:0:8:1 2>&1 | tail -4 | tee gpurun_out/lbvh_sah_dragon.log
timeout 900 python profiles/lbvh_probe.py grid448_4k_shadow 0:8:8:1.5,2:0:8:1.5,2:0:8:1 2>&1 | tail -4 | tee gpurun_out/lbvh_sah_grid.log

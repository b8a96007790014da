#!/bin/bash
# CTA-level accumulation of the bin / assign passes from the level with <= 320 members per task on, against the warp-aggregated
# passes everywhere (CERES_BUILD_WARP_PASSES=1): builder tests, build trace of the 10.5 M-triangle scene both ways.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/sah_trace_cta.log 2>&1; tail -28 gpurun_out/sah_trace_cta.log | awk '{print $3, $12, $13}' | tr '\n' ';'; tail -1 gpurun_out/sah_trace_cta.log
CERES_BUILD_WARP_PASSES=1 timeout 600 python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/sah_trace_warp.log 2>&1; tail -1 gpurun_out/sah_trace_warp.log
timeout 600 python profiles/probes/sah_build_trace.py dragon_4k_shadow 2>/dev/null | tail -1

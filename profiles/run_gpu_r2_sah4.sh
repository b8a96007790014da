#!/bin/bash
# After fusing the bottom-up box and cost climbs and warp-reducing the scene bounds: builder tests, build time of both scenes, the
# 10.5 M-triangle tree test.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python profiles/probes/sah_build_trace.py grid448_4k_shadow 2>/dev/null | tail -2 | tee gpurun_out/sah_build_fused_fit.log
timeout 600 python profiles/probes/sah_build_trace.py grid448_4k_shadow 3 2>/dev/null | tail -1
timeout 600 python profiles/probes/sah_build_trace.py dragon_4k_shadow 2>/dev/null | tail -1
timeout 900 python -m pytest tests/test_gpu_prep.py -m gpu -x -q 2>&1 | tail -3

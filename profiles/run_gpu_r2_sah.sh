#!/bin/bash
# Top-down binned SAH build (algorithm 2): builder tests, then what each tree costs at render time (dragon 4K, grid 4K).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lbvh.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/pytest_lbvh_sah.log
timeout 600 python profiles/lbvh_probe.py dragon_4k_shadow 0:8:8:1.5,2:0:8:1.5,2:0:8:1,2:0:1:1 2>&1 | tail -6 | tee gpurun_out/lbvh_sah_dragon.log
timeout 900 python profiles/lbvh_probe.py grid448_4k_shadow 0:8:8:1.5,2:0:8:1.5 2>&1 | tail -4 | tee gpurun_out/lbvh_sah_grid.log

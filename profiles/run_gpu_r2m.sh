#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "retained or host" 2>&1 | tail -5
timeout 300 python profiles/e2e_probe.py 0,4,8,12,16,24 > gpurun_out/e2e_probe_r2m.log 2>&1; cat gpurun_out/e2e_probe_r2m.log
CERES_TRACE_TIMING=1 timeout 100 python profiles/e2e_probe.py 0 2>&1 | grep "timing" | sed -n '40,42p;75,77p'

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > gpurun_out/pytest_gpu_r2o.log; cat gpurun_out/pytest_gpu_r2o.log
timeout 300 python profiles/kbench.py --share 1 --frames 20 2>&1 | tail -1
timeout 300 python profiles/e2e_probe.py 0,2,3,4,6 > gpurun_out/e2e_probe_r2o.log 2>&1; cat gpurun_out/e2e_probe_r2o.log

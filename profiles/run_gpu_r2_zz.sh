#!/bin/bash
# The static-scratch test alone in a fresh process, then the whole GPU suite (it is collected last there).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -15 > gpurun_out/pytest_zz_alone.log; cat gpurun_out/pytest_zz_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu_final4.log; cat gpurun_out/pytest_gpu_final4.log

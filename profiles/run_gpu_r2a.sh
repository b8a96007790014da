#!/bin/bash
# Round-2 session A: first run of the fused frame kernel.  Parity tests first (under timeout: a protocol bug must not hang
# the box), then the kernel A/B against the round-1 wave pipeline (experiments build) on the same box.
# Run with: gpurun --timeout 1200 -- 'bash profiles/run_gpu_r2a.sh'
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
cat gpurun_out/smoke.log | tail -5
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -40 > gpurun_out/pytest_gpu.log; echo "pytest exit ${PIPESTATUS[0]}" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python profiles/kbench.py --variants 0 --order 1,2 --frames 20 > gpurun_out/kbench_fused.log 2>&1
timeout 300 python profiles/kbench.py --variants 0 --order 1,2 --ctas 6,7 --frames 20 >> gpurun_out/kbench_fused.log 2>&1
CERES_B200_EXPERIMENTS=1 timeout 300 python profiles/kbench.py --variants 6,0 --frames 20 > gpurun_out/kbench_exp.log 2>&1
timeout 300 python profiles/kbench.py --variants 0 --order 1,2 --workload grid448_4k_shadow --frames 5 > gpurun_out/kbench_grid.log 2>&1
cat gpurun_out/kbench_fused.log gpurun_out/kbench_exp.log gpurun_out/kbench_grid.log

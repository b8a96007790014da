#!/bin/bash
# Round-1 GPU session: parity tests, smoke, bench, ncu launch list, ncu --set full of the frame's kernels.
# Run with: gpurun --timeout 1500 -- 'bash profiles/run_gpu_r1.sh'
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpu.txt
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/pytest_gpu.log; echo "pytest exit ${PIPESTATUS[0]}" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log
python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
if [ -s gpurun_out/bench.json ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:'trace_kernel|shade_kernel|clear_kernel' -s 12 -c 4 \
      -o gpurun_out/frame_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
fi
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
python bench.py --workload dragon_turntable_4k --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_seq4k_n1.json 2> gpurun_out/bench_seq4k_n1.err; tail -2 gpurun_out/bench_seq4k_n1.err
python profiles/f64_probe.py > gpurun_out/f64_probe.log 2>&1; tail -2 gpurun_out/f64_probe.log
python profiles/lbvh_probe.py grid448_4k_shadow > gpurun_out/lbvh_probe.log 2>&1; tail -7 gpurun_out/lbvh_probe.log

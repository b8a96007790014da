#!/bin/bash
# Round-2 ncu session (one GPU): launch list of the bench command, then --set full captures of the frame kernel
# (whole 4K frame with sharing off = the default at N = 1, and on; rank 0's share of 8 with sharing on and off).
# Every profiled command has first run WITHOUT ncu and exited 0.
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustain-s 0.05 > gpurun_out/bench_pre_ncu.json 2> gpurun_out/bench_pre_ncu.err || { echo "bench failed"; tail -5 gpurun_out/bench_pre_ncu.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_4k_shadow_r2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sustain-s 0.05 > gpurun_out/ncu_launches.log 2>&1; echo "launch list exit $?"
grep -c frame_kernel gpurun_out/launches_4k_shadow_r2.csv
python profiles/kbench.py --share 1,2 --frames 1 > /dev/null 2>&1 || { echo "kbench failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:frame_kernel -c 10 -o gpurun_out/frame_full_r2 -f \
    python profiles/kbench.py --share 1,2 --frames 1 > gpurun_out/ncu_full.log 2>&1; echo "full capture exit $?"
python profiles/probes/shard_alone.py --shards 8 --share 1 --frames 1 > /dev/null 2>&1 || { echo "shard_alone failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:frame_kernel -c 4 -o gpurun_out/frame_full_r2_n8_off -f \
    python profiles/probes/shard_alone.py --shards 8 --share 1 --frames 1 > gpurun_out/ncu_full_n8_off.log 2>&1; echo "n8 off exit $?"
ncu --set full --clock-control none --import-source on -k regex:frame_kernel -c 4 -o gpurun_out/frame_full_r2_n8_on -f \
    python profiles/probes/shard_alone.py --shards 8 --share 2 --frames 1 > gpurun_out/ncu_full_n8_on.log 2>&1; echo "n8 on exit $?"
ls -la gpurun_out/*.ncu-rep

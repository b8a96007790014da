#!/bin/bash
# ncu of the device tree build (N3) on the 10.5 M-triangle scene: the launch list of one build, then --set full of the level
# kernels at a mid level (a few thousand tasks) and at a deep one (> 10^6 tasks).  The command has first run without ncu.
mkdir -p gpurun_out
python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/sah_pre_ncu.log 2>&1 || { echo "probe failed"; tail -5 gpurun_out/sah_pre_ncu.log; exit 1; }
tail -1 gpurun_out/sah_pre_ncu.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_sah_build_grid.csv \
    python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/ncu_sah_launches.log 2>&1; echo "launch list exit $?"
grep -c sah_ gpurun_out/launches_sah_build_grid.csv
# first build of the probe: levels 1..27 -> launch k of sah_bin_kernel is level 6 + k (levels 1-6 use the small kernel)
ncu --set full --clock-control none --import-source on -k regex:"sah_bin_kernel|sah_assign_kernel" -s 12 -c 2 -o gpurun_out/sah_full_mid -f \
    python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/ncu_sah_mid.log 2>&1; echo "mid capture exit $?"
ncu --set full --clock-control none --import-source on -k regex:"sah_bin_kernel|sah_assign_kernel|sah_split_kernel" -s 50 -c 4 -o gpurun_out/sah_full_deep -f \
    python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/ncu_sah_deep.log 2>&1; echo "deep capture exit $?"
ls -la gpurun_out/*.ncu-rep

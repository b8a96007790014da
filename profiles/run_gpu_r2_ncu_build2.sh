#!/bin/bash
# ncu --set full of the CTA-level bin / assign passes of the tree build at a deep level (10.5 M triangles, ~1.5 M small tasks).
mkdir -p gpurun_out
python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/sah_pre_ncu.log 2>&1 || { echo "probe failed"; exit 1; }
tail -1 gpurun_out/sah_pre_ncu.log
# CTA kernels start at level 16: launch pairs (bin, assign) -> level 21 = index 10, 11
ncu --set full --clock-control none --import-source on -k regex:"sah_bin_cta_kernel|sah_assign_cta_kernel" -s 10 -c 2 -o gpurun_out/sah_full_cta_deep -f \
    python profiles/probes/sah_build_trace.py grid448_4k_shadow > gpurun_out/ncu_sah_cta.log 2>&1; echo "capture exit $?"
ls -la gpurun_out/sah_full_cta_deep.ncu-rep

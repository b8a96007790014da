#!/bin/bash
# The memcheck / racecheck of this repository.  compute-sanitizer is CLOSED on the GPU pool ("runs under it have left GPUs
# needing a reset ... find a bad access with bounds checks and asserts of your own, small cases, and a comparison with the
# CPU reference"), so the reduced workload of profiles/sanitize_driver.cpp -- every shipped kernel through every entry point
# of include/ceres_b200.h at 320x180 -- runs
#   1. against the product library (results compared call against call, 40 repeat-and-compare frames over band counts,
#      sharing on/off and grid sizes: a race shows up as a differing frame),
#   2. against the CHECKED build (-DCERES_CHECKED, csrc/checked.cuh): every data-dependent index is compared with its bound
#      on the device, and ceres_scene_frame_error verifies between frames that the queue state was re-armed,
#   3. against the checked build with CERES_CHECK_SELFTEST=1 (the triangle bound is halved): the checks MUST fire.
#   bash profiles/run_sanitizers.sh [out_dir]
# Exits non-zero if 1 or 2 report anything or 3 reports nothing.  If compute-sanitizer is usable (CERES_TRY_COMPUTE_SANITIZER=1)
# its memcheck is run as well.
ROOT=$(cd "$(dirname "$0")/.." && pwd)
OUT=${1:-$ROOT/gpurun_out}
mkdir -p "$OUT"
LIBDIR="$ROOT/ceres_raytracer_b200"
python "$ROOT/profiles/sanitize_dump.py" "$OUT" > /dev/null || exit 2
[ -f "$LIBDIR/libceres_b200_checked.so" ] || python -m ceres_raytracer_b200.build --variant checked -DCERES_CHECKED > /dev/null || exit 2
NV="nvcc -O2 -std=c++17 -Wno-deprecated-gpu-targets -I$ROOT/include"
$NV -o "$OUT/sanitize_driver" "$ROOT/profiles/sanitize_driver.cpp" -L"$LIBDIR" -lceres_b200 -Xlinker -rpath="$LIBDIR" || exit 2
$NV -o "$OUT/sanitize_driver_checked" "$ROOT/profiles/sanitize_driver.cpp" -L"$LIBDIR" -l:libceres_b200_checked.so -Xlinker -rpath="$LIBDIR" || exit 2
bad=0
timeout 600 "$OUT/sanitize_driver" "$OUT/scene.bin" "$OUT/scene64.bin" > "$OUT/sanitizer_product.log" 2>&1; rc=$?
echo "product library: exit $rc; $(tail -1 "$OUT/sanitizer_product.log")"; [ $rc -ne 0 ] && bad=1
timeout 600 "$OUT/sanitize_driver_checked" "$OUT/scene.bin" "$OUT/scene64.bin" > "$OUT/sanitizer_checked.log" 2>&1; rc=$?
echo "checked build:   exit $rc; $(tail -1 "$OUT/sanitizer_checked.log")"; [ $rc -ne 0 ] && bad=1
grep -q "CHECKED build" "$OUT/sanitizer_checked.log" || { echo "checked build: the driver did not load the checked library"; bad=1; }
CERES_CHECK_SELFTEST=1 timeout 600 "$OUT/sanitize_driver_checked" "$OUT/scene.bin" "$OUT/scene64.bin" > "$OUT/sanitizer_selftest.log" 2>&1; rc=$?
n=$(grep -c "DEVICE CHECK FAILED\|an index left its bound\|error bits\|differs" "$OUT/sanitizer_selftest.log")
echo "self-test (bound halved): exit $rc, $n report(s) -- $( [ "$n" -gt 0 ] && echo 'the checks are live' || echo 'NOTHING REPORTED: the checks are dead' )"
[ "$n" -gt 0 ] || bad=1
if [ -n "$CERES_TRY_COMPUTE_SANITIZER" ]; then
  CS=$(command -v compute-sanitizer || echo /usr/local/cuda/bin/compute-sanitizer)
  timeout 1500 "$CS" --tool memcheck --error-exitcode 77 "$OUT/sanitize_driver" "$OUT/scene.bin" "$OUT/scene64.bin" > "$OUT/sanitizer_memcheck.log" 2>&1
  echo "compute-sanitizer memcheck: exit $?; $(grep 'ERROR SUMMARY' "$OUT/sanitizer_memcheck.log" || head -1 "$OUT/sanitizer_memcheck.log")"
fi
exit $bad

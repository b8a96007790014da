# e2e of the 4K host frame vs the size / store flavour of the host fill pool (alternating runs on one box)
for v in "" "CERES_HOST_FILL_THREADS=8" "CERES_HOST_FILL_THREADS=11" "CERES_HOST_FILL_THREADS=15" "" "CERES_HOST_FILL_THREADS=8" "CERES_HOST_FILL_THREADS=11" "CERES_HOST_FILL_CACHED=1"; do
  echo "== ${v:-default (7 pool threads)}"; env $v python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('e2e ms', round(d['e2e']['ms_per_step'],4), 'ppm8', round(d['e2e_ppm8']['ms_per_step'],4), 'ppm16', round(d['e2e_ppm8']['ppm16']['ms_per_step'],4))"
done

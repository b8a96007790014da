# A/B of the tight band copies on one box: default vs CERES_NO_TIGHT_COPY=1 (conservative-rectangle copies enqueued behind each band)
for v in "" "CERES_NO_TIGHT_COPY=1" "" "CERES_NO_TIGHT_COPY=1"; do
  echo "== ${v:-tight (default)}"; env $v python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('e2e ms', round(d['e2e']['ms_per_step'],4), 'd2h bytes', d['e2e']['d2h_bytes_per_step'], 'frame ms', round(d['ms_per_step'],4))"
done

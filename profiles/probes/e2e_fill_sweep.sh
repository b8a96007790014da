for v in "" "CERES_HOST_FILL_CACHED=1" "CERES_HOST_FILL_THREADS=15" "CERES_HOST_FILL_THREADS=5" "CERES_HOST_FILL_THREADS=0"; do
  echo "== $v"; env $v python bench.py --steps 30 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['e2e']['ms_per_step'], d['e2e_ppm8']['ms_per_step'], d['ms_per_step'])"
done

for mode in "" 1; do
  export CERES_SEQUENCE_SERIAL=$mode; [ -z "$mode" ] && unset CERES_SEQUENCE_SERIAL
  timeout 600 python bench.py --workload dragon_turntable_4k --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('serial' if '$mode' else 'two in flight', 'e2e ms/frame', round(d['e2e']['ms_per_step']/60,3), 'ppm8', round(d['e2e_ppm8']['ms_per_step']/60,3))"
done

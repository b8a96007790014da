#!/bin/bash
# All BASELINE.json configurations on one B200, each with --verify (the same frame rendered by the CPU oracle at full
# size: hit ids, counters, RGB compared).  Run with: gpurun --timeout 1500 -- 'bash profiles/run_gpu_configs.sh'
mkdir -p gpurun_out
for w in bunny_1080p_primary bunny_1080p_shadow dragon_1080p_primary dragon_1080p_shadow dragon_4k_shadow grid448_8k_shadow; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --verify > gpurun_out/cfg_$w.json 2> gpurun_out/cfg_$w.err
  echo "$w exit $?"
  python -c "
import json,sys; d=json.load(open('gpurun_out/cfg_$w.json')); print(round(d['value']), round(d['ms_per_step'],3), round(d['e2e']['value']), round(d['e2e']['ms_per_step'],3), d.get('verified'))"
done

#!/bin/bash
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
for n in 8 4; do
  [ "$n" -le "$NG" ] || continue
  for sh in 1 2; do
    SHARE=$sh timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2953$n profiles/probes/peer_timeline.py 2>&1 | grep -v Warning | tail -24 >> gpurun_out/peer_timeline.log
  done
done
cat gpurun_out/peer_timeline.log

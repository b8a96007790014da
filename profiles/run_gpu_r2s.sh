#!/bin/bash
for v in "" ctas7 ctas6 ctas10; do
  echo "== variant '$v'"
  CERES_B200_VARIANT=$v timeout 300 python profiles/kbench.py --share 1 --frames 20 2>&1 | tail -1
done

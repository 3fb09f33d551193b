#!/bin/bash
TAG=${1:-j7}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for st in 0 200 500 1000 2000 4000; do
HPCX_TB_STAGGER=$st TB=1 ZS=0 timeout 300 python scripts/prof_j7.py 2>&1 | sed "s/^/stagger$st /" | tee -a $OUT/tb.txt
done

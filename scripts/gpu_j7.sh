#!/bin/bash
TAG=${1:-j7}
OUT=gpurun_out/$TAG
mkdir -p $OUT
TB=1 ZS=0 timeout 300 python scripts/prof_j7.py 2>&1 | tee -a $OUT/tb.txt
HPCX_TB_FORCE_PEER=1 TB=1 ZS=0 timeout 300 python scripts/prof_j7.py 2>&1 | sed "s/^/forcepeer /" | tee -a $OUT/tb.txt

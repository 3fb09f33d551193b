#!/bin/bash
# Short GPU pass: parity tests + bench (no ncu).  Usage: bash scripts/gpu_quick.sh <tag> [pytest args]
TAG=${1:-q}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -15 $OUT/pytest_gpu.log
timeout 600 python bench.py --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 2500 $OUT/bench.json; tail -3 $OUT/bench.err

#!/bin/bash
# Short GPU pass: parity tests + bench (no ncu).  Usage: bash scripts/gpu_quick.sh <tag> [pytest args]
TAG=${1:-q}
shift
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q "$@" > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -15 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 3500 $OUT/bench.json; tail -5 $OUT/bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" | tee -a $OUT/rc.txt
tail -c 600 $OUT/bench_ref.json

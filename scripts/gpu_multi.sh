#!/bin/bash
# Multi-GPU pass: distributed parity tests + bench at N GPUs.  Usage: bash scripts/gpu_multi.sh <tag> <N>
TAG=${1:-m}
N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > $OUT/pytest_dist.log 2>&1; echo "pytest-dist rc=$?" | tee $OUT/rc.txt
tail -25 $OUT/pytest_dist.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --no-cpu > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 3000 $OUT/bench_n$N.json; tail -5 $OUT/bench_n$N.err

#!/bin/bash
# GPU pass for the C4/C5 extras: parity tests, bench (N=1, all configs), ncu --set full of the fp32 (R-MAT) and
# fp64 27-point instances of the stream kernel.  Usage: bash scripts/gpu_c45.sh <tag>
TAG=${1:-c45}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -5 $OUT/pytest_gpu.log
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 2500 $OUT/bench.json; tail -5 $OUT/bench.err
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
  -k 'regex:spmv_stream_kernel<float' --launch-skip 3 -c 1 -o $OUT/spmv_stream_rmat_full -f \
  python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/ncu_rmat.log 2>&1; echo "ncu-rmat rc=$?" | tee -a $OUT/rc.txt
fi

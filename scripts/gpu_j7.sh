#!/bin/bash
TAG=${1:-j7}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_stencil.py tests/test_gpu_dist.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -5 $OUT/pytest.log
for n in 256 384 512 640 768; do
N=$n ZS=0 timeout 300 python scripts/prof_j7.py 2>&1 | tee -a $OUT/auto.txt
N=$n TB=1 ZS=0 timeout 300 python scripts/prof_j7.py 2>&1 | tee -a $OUT/auto.txt
N=$n TB=1 HPCX_TB_SLOTS=5 ZS=0,3,6,9 timeout 300 python scripts/prof_j7.py 2>&1 | sed "s/^/slots5 /" | tee -a $OUT/auto.txt
done
HPCX_TB_SLOTS=5 timeout 600 python -m pytest tests/test_gpu_stencil.py -m gpu -x -q -k "temporal" > $OUT/pytest5.log 2>&1; echo "pytest5 rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest5.log

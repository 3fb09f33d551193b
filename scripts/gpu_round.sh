#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (N=1), ncu launch list, ncu --set full of the top kernels.
# Usage (from the repo root, under gpurun): bash scripts/gpu_round.sh [tag]
TAG=${1:-r01}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $OUT/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q --timeout=600 --timeout-method=thread > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
timeout 300 python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
tail -c 3000 $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 5 --warmup 2 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?" | tee -a $OUT/rc.txt
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/launches.csv \
  python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/ncu_launch.log 2>&1; echo "ncu-launches rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:spmv_stream_kernel<double" --launch-skip 4 -c 1 \
  -o $OUT/spmv_stream_full -f python bench.py --steps 3 --warmup 3 --no-cpu --configs headline > $OUT/ncu_spmv.log 2>&1; echo "ncu-spmv rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:jacobi7_ring_kernel --launch-skip 10 -c 1 \
  -o $OUT/jacobi7_ring_full -f python bench.py --steps 3 --warmup 3 --no-cpu --configs headline > $OUT/ncu_jacobi.log 2>&1; echo "ncu-jacobi rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:jacobi7_tb2_kernel --launch-skip 10 -c 1 \
  -o $OUT/jacobi7_tb2_full -f python bench.py --steps 3 --warmup 3 --no-cpu --configs headline > $OUT/ncu_tb2.log 2>&1; echo "ncu-tb2 rc=$?" | tee -a $OUT/rc.txt
# C5 (R-MAT scale 24): the HUB kernel, captured from the check script (20 timed launches: skip into them)
FULL=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmv_hub_kernel --launch-skip 20 -c 1 \
  -o $OUT/spmv_hub_rmat_full -f python scripts/hub_check.py > $OUT/ncu_hub.log 2>&1; echo "ncu-hub rc=$?" | tee -a $OUT/rc.txt
fi
cat $OUT/rc.txt

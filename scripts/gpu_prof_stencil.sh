#!/bin/bash
TAG=${1:-p}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_stencil.py -m gpu -x -q > $OUT/pytest_stencil.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -3 $OUT/pytest_stencil.log
timeout 300 python scripts/prof_stencil.py all 20 > $OUT/stencil_times.txt 2>&1 && cat $OUT/stencil_times.txt &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'jacobi7_tma_kernel|box_y_reg2_kernel|box_x_vec_kernel' --launch-skip 6 -c 1 \
  -o $OUT/jacobi7_tma_full -f python scripts/prof_stencil.py jacobi 3 > $OUT/ncu_jacobi.log 2>&1; echo "ncu-jacobi rc=$?" | tee -a $OUT/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'box_y_reg2_kernel' --launch-skip 2 -c 1 \
  -o $OUT/box_y_full -f python scripts/prof_stencil.py box 3 > $OUT/ncu_boxy.log 2>&1; echo "ncu-boxy rc=$?" | tee -a $OUT/rc.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'box_x_vec_kernel' --launch-skip 2 -c 1 \
  -o $OUT/box_x_full -f python scripts/prof_stencil.py box 3 > $OUT/ncu_boxx.log 2>&1; echo "ncu-boxx rc=$?" | tee -a $OUT/rc.txt

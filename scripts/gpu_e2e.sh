#!/bin/bash
TAG=${1:-e2e}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_spmv.py tests/test_cpp_host.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee $OUT/rc.txt
tail -8 $OUT/pytest.log
timeout 600 python bench.py --no-cpu --configs headline > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
d=json.loads(open("$OUT/bench.json").read().strip().splitlines()[-1])
print(d["value"], d["e2e"])
PY

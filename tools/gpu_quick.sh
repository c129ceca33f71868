#!/bin/bash
# tools/gpu_quick.sh -- one short gpurun call while iterating on the MPC kernel: MPC parity tests, then the bench line
# without the auxiliary legs.  usage: gpurun --timeout 900 -- 'bash tools/gpu_quick.sh <tag>'
set -u
TAG=${1:-quick}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_mpc_parity.py tests/test_gpu_golden.py -m gpu -x -q > $OUT/pytest_mpc.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest_mpc.log
timeout 600 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --wbc-batch 0 > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"], 3), "iters", d["admm_iters"], "clocks", d["clocks"])
except Exception as e:
    print("no bench line:", e)
    print(open("$OUT/bench.err").read()[-2000:])
PY

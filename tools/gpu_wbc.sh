#!/bin/bash
# tools/gpu_wbc.sh -- whole-body-tick tests and the bench line with the config-4 leg (65536 robots).
set -u
TAG=${1:-wbc}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_wbc_tick.py tests/test_gpu_wbc_invkin.py tests/test_gpu_controller.py tests/test_gpu_tick_abi.py tests/test_gpu_golden.py -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -15 $OUT/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --steps 4 --warmup 3 --no-cpu-baseline > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench.json").read().strip().splitlines()[-1])
    print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]))
    for k in ("wbc_tick", "full_tick", "planner_tick", "control_loop"):
        print(k, round(d[k]["value"]), d[k].get("ms_per_step"))
except Exception as e:
    print("no bench line:", e)
    print(open("$OUT/bench.err").read()[-2000:])
PY

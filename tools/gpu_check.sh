#!/bin/bash
# tools/gpu_check.sh -- one gpurun call: GPU parity tests, bench line, ncu launch list, one full ncu capture.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh <tag>'
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -5 $OUT/pytest_gpu.log
timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
cat $OUT/bench.json
SMALL="python bench.py --batch 592 --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $SMALL > $OUT/small.json 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches.csv $SMALL > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
timeout 300 $SMALL > $OUT/small2.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_v2_kernel -s 6 -c 1 -f -o $OUT/prof_mpc $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT

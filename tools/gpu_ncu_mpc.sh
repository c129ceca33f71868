#!/bin/bash
# tools/gpu_ncu_mpc.sh -- one full ncu capture of the default MPC kernel on the bench workload (4096 robots, one update tick).
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_ncu_mpc.sh <tag> [kernel regex]'
set -u
TAG=${1:-ncu}
K=${2:-mpc_v3_kernel}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SMALL="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --wbc-batch 0"
timeout 300 $SMALL > $OUT/small.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s 5 -c 1 -f -o $OUT/prof_mpc $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu mpc rc=$?" | tee -a $OUT/rc.txt
ncu -i $OUT/prof_mpc.ncu-rep --page details > $OUT/prof_mpc_details.txt 2>/dev/null
tail -3 $OUT/small.json | cut -c1-300

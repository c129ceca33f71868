#!/bin/bash
# tools/gpu_profile_planner.sh -- ncu --set full of planner_tick_kernel (65,536 robots; one MPC tick k = 10 and one plain
# tick k = 11) after the same command exited 0 without ncu.  usage: gpurun --timeout 900 -- 'bash tools/gpu_profile_planner.sh <tag>'
set -u
TAG=${1:-plan}
OUT=gpurun_out/$TAG
mkdir -p $OUT
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $OUT/plain.json 2> $OUT/plain.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:planner_tick_kernel -s 10 -c 2 -f -o $OUT/prof_planner $CMD > $OUT/ncu.log 2>&1
echo "ncu planner rc=$?"
ncu -i $OUT/prof_planner.ncu-rep --page raw --csv > $OUT/prof_planner_raw.csv 2>/dev/null
ls -la $OUT

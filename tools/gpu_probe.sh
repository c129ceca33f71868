#!/bin/bash
# tools/gpu_probe.sh -- timing probe (clock64 build) + one full ncu capture of mpc_v2_kernel.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_probe.sh <tag>'
set -u
TAG=${1:-probe}
OUT=gpurun_out/$TAG
mkdir -p $OUT
[ -f tools/libqrw_timing.so ] && QRW_LIB=$PWD/tools/libqrw_timing.so timeout 300 python tools/timing_probe.py > $OUT/timing_probe.txt 2>&1
echo "probe rc=$?" | tee -a $OUT/rc.txt
cat $OUT/timing_probe.txt
SMALL="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --wbc-batch 0"
timeout 300 $SMALL > $OUT/small.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_v2_kernel -s 5 -c 1 -f -o $OUT/prof_mpc $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu mpc rc=$?" | tee -a $OUT/rc.txt
ncu -i $OUT/prof_mpc.ncu-rep --page raw --csv > $OUT/prof_mpc_raw.csv 2>/dev/null
ncu -i $OUT/prof_mpc.ncu-rep --page details > $OUT/prof_mpc_details.txt 2>/dev/null
ls -la $OUT

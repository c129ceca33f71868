#!/bin/bash
# tools/gpu_check.sh -- one gpurun call: GPU parity tests, smoke, bench line, ncu launch list, full ncu captures.
# usage: gpurun --timeout 1800 -- 'bash tools/gpu_check.sh <tag>'
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
cat $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2>> $OUT/bench.err; echo "bench reference rc=$?" | tee -a $OUT/rc.txt
LIST="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $LIST > $OUT/list_plain.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $LIST > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
SMALL="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --wbc-batch 4096"
timeout 300 $SMALL > $OUT/small.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_v2_kernel -s 5 -c 1 -f -o $OUT/prof_mpc $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu mpc rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wbc_tick_pre_kernel|wbc_solve_kernel|wbc_tick_post_kernel" -s 9 -c 3 -f -o $OUT/prof_wbc $SMALL > $OUT/ncu_wbc.log 2>&1
echo "ncu wbc rc=$?" | tee -a $OUT/rc.txt
ls -la $OUT
timeout 900 ncu --set full --clock-control none --import-source on -k regex:planner_tick_kernel -s 10 -c 2 -f -o $OUT/prof_planner $SMALL > $OUT/ncu_planner.log 2>&1
echo "ncu planner rc=$?" | tee -a $OUT/rc.txt
for r in prof_mpc prof_wbc prof_planner; do ncu -i $OUT/$r.ncu-rep --page raw --csv > $OUT/${r}_raw.csv 2>/dev/null; done
ls -la $OUT
timeout 60 tools/chain_bench.bin > $OUT/chain_bench.txt 2>&1; timeout 60 tools/chain_bench_t.bin >> $OUT/chain_bench.txt 2>&1
[ -f tools/libqrw_timing.so ] && QRW_LIB=$PWD/tools/libqrw_timing.so timeout 200 python tools/timing_probe.py > $OUT/timing_probe.txt 2>&1
echo "probes rc=$?" | tee -a $OUT/rc.txt

#!/bin/bash
# tools/gpu_check.sh -- one gpurun call: GPU parity tests, smoke, the bench lines (default config 2, reference arm, configs 3
# and 5 on one GPU), ncu launch list, full ncu captures.  usage: gpurun --timeout 2400 -- 'bash tools/gpu_check.sh <tag> [noncu]'
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/smi.txt 2>&1
nproc > $OUT/nproc.txt; free -g >> $OUT/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/rc.txt
tail -3 $OUT/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py --impl reference --steps 8 --warmup 3 > $OUT/bench_reference.json 2> $OUT/bench.err; echo "bench reference rc=$?" | tee -a $OUT/rc.txt
timeout 900 python bench.py > $OUT/bench.json 2>> $OUT/bench.err; echo "bench rc=$?" | tee -a $OUT/rc.txt
cat $OUT/bench.json | cut -c1-1500
timeout 900 python bench.py --config 3 --steps 4 --warmup 3 > $OUT/bench_config3.json 2>> $OUT/bench.err; echo "bench config3 rc=$?" | tee -a $OUT/rc.txt
cut -c1-700 $OUT/bench_config3.json
timeout 1500 python bench.py --config 5 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/bench_config5_1gpu.json 2>> $OUT/bench.err; echo "bench config5 rc=$?" | tee -a $OUT/rc.txt
cut -c1-700 $OUT/bench_config5_1gpu.json
[ "${2:-}" = "noncu" ] && exit 0
LIST="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 $LIST > $OUT/list_plain.json 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $LIST > $OUT/ncu_launches.log 2>&1
echo "ncu launches rc=$?" | tee -a $OUT/rc.txt
SMALL="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --wbc-batch 65536"
timeout 300 $SMALL > $OUT/small.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_v2_kernel -s 5 -c 1 -f -o $OUT/prof_mpc $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu mpc rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wbc_tick_kernel" -s 4 -c 2 -f -o $OUT/prof_wbc $SMALL > $OUT/ncu_wbc.log 2>&1
echo "ncu wbc rc=$?" | tee -a $OUT/rc.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:planner_tick_kernel -s 10 -c 2 -f -o $OUT/prof_planner $SMALL > $OUT/ncu_planner.log 2>&1
echo "ncu planner rc=$?" | tee -a $OUT/rc.txt
for r in prof_mpc prof_wbc prof_planner; do ncu -i $OUT/$r.ncu-rep --page details > $OUT/${r}_details.txt 2>/dev/null; done
ls -la $OUT

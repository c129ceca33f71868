#!/bin/bash
# tools/gpu_multi.sh -- N-GPU bench lines: default config (weak scaling) and config 5 (1,048,576 robots, strong scaling).
# usage: gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_multi.sh <tag> N'
set -u
TAG=${1:-multi}
N=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $RUN bench.py --gpus $N --steps 8 --warmup 3 --no-cpu-baseline > $OUT/bench_${N}gpu.json 2> $OUT/bench_${N}gpu.err; echo "bench $N rc=$?" | tee -a $OUT/rc.txt
tail -1 $OUT/bench_${N}gpu.json | cut -c1-400
timeout 1200 $RUN bench.py --gpus $N --config 5 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/bench_config5_${N}gpu.json 2> $OUT/bench_config5_${N}gpu.err; echo "bench config5 $N rc=$?" | tee -a $OUT/rc.txt
tail -1 $OUT/bench_config5_${N}gpu.json | cut -c1-400

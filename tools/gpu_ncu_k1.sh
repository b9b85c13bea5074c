#!/bin/bash
# ncu --set full (+ source) of k1_search, after the same command exited 0 without ncu; usage: gpu_ncu_k1.sh <out name> [kernel regex]
set -x
mkdir -p gpurun_out
NAME=${1:-k1}; KRE=${2:-k1_search}
CMD="python bench.py --nkf 24 --steps 1 --warmup 1 --no-cpu-baseline --no-configs"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 30 -c 2 -f -o gpurun_out/$NAME $CMD > gpurun_out/ncu_$NAME.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$NAME.log

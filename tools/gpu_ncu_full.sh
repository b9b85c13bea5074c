#!/bin/bash
# ncu --set full of the two heaviest kernels (after the same command exited 0 without ncu)
set -x
mkdir -p gpurun_out
CMD="python bench.py --nkf 24 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k1_search -s 30 -c 2 -f -o gpurun_out/k1_full $CMD > gpurun_out/ncu_k1.log 2>&1
echo "ncu k1 rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k4_inter -s 30 -c 2 -f -o gpurun_out/k4_full $CMD > gpurun_out/ncu_k4.log 2>&1
echo "ncu k4 rc=$?"

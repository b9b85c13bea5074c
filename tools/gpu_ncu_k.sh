#!/bin/bash
# ncu --set full of kernels: usage gpu_ncu_k.sh <regex> <output name> [skip] ; repeatable as triples
mkdir -p gpurun_out
CMD="python bench.py --nkf 24 --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
while [ $# -ge 2 ]; do
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$1 -s ${3:-30} -c 1 -f -o gpurun_out/$2 $CMD > gpurun_out/ncu_$2.log 2>&1
  echo "ncu $2 rc=$?"
  shift 3
done

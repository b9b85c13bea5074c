#!/bin/bash
# bench value for several settings of one environment variable: usage gpu_sweep_env.sh VAR v1 v2 ...
var=$1; shift
mkdir -p gpurun_out
for b in "$@"; do
  echo "=== $var=$b"
  env $var=$b timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | tee -a gpurun_out/sweep_$var.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value %.1f e2e %.1f  kernels_avg_us %s' % (d['value'], d['e2e']['value'], {k: round(v,1) for k,v in d['kernels_avg_us'].items()}))"
done

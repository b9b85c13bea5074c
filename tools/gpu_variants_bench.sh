#!/bin/bash
# build variants of libsdpm.so on the GPU box and bench each: usage gpu_variants_bench.sh "<nvcc flags>|<ENV=V ...>" ...
mkdir -p gpurun_out
for v in "$@"; do
  flags="${v%%|*}"; envs="${v#*|}"; [ "$envs" = "$v" ] && envs=""
  echo "=== flags: $flags   env: $envs"
  SDPM_NVCC_EXTRA="$flags" python orb-slam-free-space-carving_b200/build.py --force > /dev/null 2>&1 || { echo build failed; continue; }
  grep -A2 "k1_search" orb-slam-free-space-carving_b200/lib/libsdpm.build.log | grep registers
  env $envs timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | tee -a gpurun_out/variants.jsonl | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value %.1f e2e %.1f  kernels_avg_us %s' % (d['value'], d['e2e']['value'], {k: round(v,1) for k,v in d['kernels_avg_us'].items()}))"
done

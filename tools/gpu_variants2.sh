#!/bin/bash
# build variants of libsdpm.so on the GPU box and bench each: usage gpu_variants2.sh "<nvcc flags>|<ENV=V ...>" ...
mkdir -p gpurun_out
for v in "$@"; do
  flags="${v%%|*}"; envs="${v#*|}"; [ "$envs" = "$v" ] && envs=""
  echo "=== flags: $flags   env: $envs"
  SDPM_NVCC_EXTRA="$flags" python orb-slam-free-space-carving_b200/build.py --force > /dev/null 2>&1 || { echo build failed; continue; }
  grep -A3 "k1_search\|k0_gradient" orb-slam-free-space-carving_b200/lib/libsdpm.build.log | grep -E "Used|spill" | head -4
  env $envs timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs 2>/dev/null | tee -a gpurun_out/variants.jsonl > gpurun_out/v.json; python tools/show_bench.py gpurun_out/v.json
done

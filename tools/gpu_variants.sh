#!/bin/bash
# build variants of libsdpm.so on the GPU box (nvcc is in the image) and time each: usage gpu_variants.sh "<flags A>" "<flags B>" ...
mkdir -p gpurun_out
for v in "$@"; do
  echo "=== variant: $v"
  SDPM_NVCC_EXTRA="$v" python orb-slam-free-space-carving_b200/build.py --force > /dev/null 2>&1 || { echo build failed; continue; }
  grep -A2 "k1_search" orb-slam-free-space-carving_b200/lib/libsdpm.build.log | grep registers
  timeout 300 python tools/gpu_check.py 720p 24 10 --no-oracle 2>&1 | grep "rep2\|KF/s"
done

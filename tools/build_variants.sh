#!/bin/bash
# Cross-compile named variants of libsdpm.so HERE (no GPU needed) so that the GPU box only has to run them:
#   tools/build_variants.sh name1 "<nvcc -D flags>" name2 "<flags>" ...
# -> orb-slam-free-space-carving_b200/lib/variants/<name>/libsdpm.so   (lib/ is git-ignored, it travels with gpurun)
set -e
cd "$(dirname "$0")/.."
PKG=orb-slam-free-space-carving_b200
while [ $# -ge 2 ]; do
  name="$1"; flags="$2"; shift 2
  mkdir -p $PKG/lib/variants/$name
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
       -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math -Xptxas -v $flags -shared -o $PKG/lib/variants/$name/libsdpm.so $PKG/csrc/sdpm_api.cu \
       > $PKG/lib/variants/$name/build.log 2>&1
  echo "$name: $flags"; grep -A2 "k1_search\|k0_gradient" $PKG/lib/variants/$name/build.log | grep "registers" | sed 's/ptxas info    ://'
done

#!/bin/bash
mkdir -p gpurun_out
for cfg in "16 4" "16 8" "32 8" "32 12" "64 8"; do
set -- $cfg
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --class-window $1 --class-threads $2 > gpurun_out/r03h_bench_w$1_t$2.json 2> gpurun_out/r03h_bench_w$1_t$2.err; echo "window $1 threads $2 rc=$?"
python tools/show_bench.py gpurun_out/r03h_bench_w$1_t$2.json | tail -1 | cut -c1-80
grep "last pass" gpurun_out/r03h_bench_w$1_t$2.err
done

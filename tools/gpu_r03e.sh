#!/bin/bash
mkdir -p gpurun_out
for w in 16 64; do
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --class-window $w > gpurun_out/r03e_bench_w$w.json 2> gpurun_out/r03e_bench_w$w.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/r03e_bench_w$w.json | tail -1
grep "last pass" gpurun_out/r03e_bench_w$w.err
done

#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_host_class.py -m gpu -x -q > gpurun_out/r03g_host.log 2>&1; echo "host rc=$?" >> gpurun_out/r03g_host.log
tail -4 gpurun_out/r03g_host.log
for w in 8 16 32; do
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --class-window $w > gpurun_out/r03g_bench_w$w.json 2> gpurun_out/r03g_bench_w$w.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/r03g_bench_w$w.json | tail -1
grep "last pass" gpurun_out/r03g_bench_w$w.err
done

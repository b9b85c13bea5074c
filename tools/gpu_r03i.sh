#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_host_class.py -m gpu -x -q > gpurun_out/r03i_host.log 2>&1; echo "host rc=$?" >> gpurun_out/r03i_host.log
tail -3 gpurun_out/r03i_host.log
for cfg in "16 0" "16 12" "32 0" "32 12" "8 0"; do
set -- $cfg
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --class-window $1 --class-threads $2 > gpurun_out/r03i_bench_w$1_t$2.json 2> gpurun_out/r03i_bench_w$1_t$2.err; echo "window $1 threads $2 rc=$?"
python tools/show_bench.py gpurun_out/r03i_bench_w$1_t$2.json | tail -1 | cut -c1-80
grep "last pass" gpurun_out/r03i_bench_w$1_t$2.err
done

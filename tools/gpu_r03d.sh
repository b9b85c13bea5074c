#!/bin/bash
# round 3: bench with the class leg; sweep of the class window / threads on the same scene
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r03d_bench.json 2> gpurun_out/r03d_bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/r03d_bench.json | head -12
tail -3 gpurun_out/r03d_bench.err
for w in 8 32 64; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-configs --class-window $w > gpurun_out/r03d_bench_w$w.json 2>/dev/null
  python -c "import json,sys; d=json.load(open('gpurun_out/r03d_bench_w$w.json')); print('window $w', d['e2e_class'].get('value'), d['e2e_class'].get('ms_per_step'), d['e2e_class'].get('unavailable'))"
done
nproc; free -g | head -2

#!/bin/bash
# quick iteration: core parity tests, then the bench legs only (no CPU baseline, no other configs)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_round2.py -m gpu -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_quick.log
tail -5 gpurun_out/pytest_quick.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/bench_quick.json 2>/dev/null || cat gpurun_out/bench_quick.json
tail -3 gpurun_out/bench_quick.err

#!/bin/bash
# round 3: whole GPU suite (-x like the driver), then a short bench
mkdir -p gpurun_out
T=${1:-r03b}
timeout 1800 python -m pytest tests -m gpu -x -q -v > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "suite rc=$?" >> gpurun_out/${T}_pytest_gpu.log
grep -v PASSED gpurun_out/${T}_pytest_gpu.log | tail -25
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/${T}_bench.json 2>/dev/null | head -12

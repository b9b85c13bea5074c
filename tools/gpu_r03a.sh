#!/bin/bash
# round 3: the no-allocation entry point alone in a fresh process, then the whole GPU suite (it is collected last), then smoke + bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -v > gpurun_out/r03a_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/r03a_noalloc_alone.log
tail -8 gpurun_out/r03a_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q -v > gpurun_out/r03a_pytest_gpu.log 2>&1; echo "suite rc=$?" >> gpurun_out/r03a_pytest_gpu.log
tail -12 gpurun_out/r03a_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r03a_smoke.log 2>&1; tail -2 gpurun_out/r03a_smoke.log
timeout 900 python bench.py > gpurun_out/r03a_bench.json 2> gpurun_out/r03a_bench.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/r03a_bench.json 2>/dev/null | head -40

#!/bin/bash
# round 3: host-class tests (windowed SemiDenseLoop), then the rest of the suite
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_host_class.py tests/test_abi.py -x -q -v > gpurun_out/r03c_host.log 2>&1; echo "host rc=$?" >> gpurun_out/r03c_host.log
grep -v PASSED gpurun_out/r03c_host.log | tail -40

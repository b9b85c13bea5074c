#!/bin/bash
# GPU parity suite only (no -x: every failure in one call)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -60 gpurun_out/pytest_gpu.log

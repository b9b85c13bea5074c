#!/bin/bash
# two GPUs: sharded parity tests (NCCL through the C ABI), bench at N=2 (per-rank spans, clocks, compute-stream timeline,
# seam parity) and at N=1 on the same box; optional prebuilt variants on one GPU
mkdir -p gpurun_out
TAG=${TAG:-r02e}
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/${TAG}_gpus.txt; nproc >> gpurun_out/${TAG}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -q -v > gpurun_out/${TAG}_pytest_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_2gpu.log
tail -6 gpurun_out/${TAG}_pytest_2gpu.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_2gpu.json 2> gpurun_out/${TAG}_bench_2gpu.err; echo "bench2 rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_2gpu.json; tail -5 gpurun_out/${TAG}_bench_2gpu.err
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/${TAG}_bench_1gpu_same_box.json 2> gpurun_out/${TAG}_bench_1gpu.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_1gpu_same_box.json
[ -d orb-slam-free-space-carving_b200/lib/variants ] && bash tools/gpu_variants_run.sh 2>&1 | tee gpurun_out/${TAG}_variants.txt

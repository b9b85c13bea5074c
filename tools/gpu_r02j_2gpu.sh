#!/bin/bash
# two GPUs: the whole GPU suite with test names (the two-rank NCCL test is not skipped here), then BASELINE config 3's
# shape sharded as stated (720p, N=20, key-frame-sharded, NCCL map exchange) with seam parity against the oracle
mkdir -p gpurun_out
TAG=${TAG:-r02j}
timeout 300 python -m pytest tests -m gpu -x -rs -v > gpurun_out/${TAG}_pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu_2gpu.log
tail -3 gpurun_out/${TAG}_pytest_gpu_2gpu.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 2 --nbr 20 --nkf ${NKF:-100} --steps 3 --warmup 3 > gpurun_out/${TAG}_bench_config3_2gpu.json 2> gpurun_out/${TAG}_bench_config3_2gpu.err; echo "bench rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_config3_2gpu.json 2>/dev/null | head -8; tail -3 gpurun_out/${TAG}_bench_config3_2gpu.err

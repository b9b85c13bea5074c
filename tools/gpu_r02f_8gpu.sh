#!/bin/bash
# one 8-GPU box: bench (all legs, per-rank spans / clocks, timeline, seam parity) at $NS (default 8 and 2) and 1 GPUs; REF_ARM=1 adds the reference arm at 8
mkdir -p gpurun_out
TAG=${TAG:-r02f}
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/${TAG}_gpus.txt; nproc >> gpurun_out/${TAG}_gpus.txt
for n in ${NS:-8 2}; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_${n}gpu.json 2> gpurun_out/${TAG}_bench_${n}gpu.err; echo "bench$n rc=$?"
  python tools/show_bench.py gpurun_out/${TAG}_bench_${n}gpu.json; tail -2 gpurun_out/${TAG}_bench_${n}gpu.err
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/${TAG}_bench_1gpu_same_box.json 2> gpurun_out/${TAG}_bench_1gpu.err; python tools/show_bench.py gpurun_out/${TAG}_bench_1gpu_same_box.json
[ -z "$REF_ARM" ] || { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29790 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/${TAG}_reference_arm_8gpu.json 2>/dev/null; cut -c1-500 gpurun_out/${TAG}_reference_arm_8gpu.json; }

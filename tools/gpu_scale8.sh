#!/bin/bash
# one 8-GPU box: D2H ceiling at 1/2/4/8 ranks, then bench (all legs + seam parity) at 8, 4, 2, 1 and the reference arm at 8
mkdir -p gpurun_out
: > gpurun_out/d2h_ceiling.jsonl
for n in 1 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) tools/d2h_ceiling.py 100 2>/dev/null >> gpurun_out/d2h_ceiling.jsonl
done
cat gpurun_out/d2h_ceiling.jsonl
for n in 8 4 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err; echo "bench$n rc=$?"
  python tools/show_bench.py gpurun_out/bench_${n}gpu.json; tail -3 gpurun_out/bench_${n}gpu.err
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/bench_1gpu_same_box.json 2> gpurun_out/bench_1gpu.err; python tools/show_bench.py gpurun_out/bench_1gpu_same_box.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29790 bench.py --impl reference --gpus 8 --steps 1 --warmup 1 > gpurun_out/ref_8gpu.json 2>/dev/null; cut -c1-700 gpurun_out/ref_8gpu.json

#!/bin/bash
# round 3, two GPUs: the sharded test (not skipped here), then the N=2 bench line with the configs block and the CPU leg
mkdir -p gpurun_out
SECONDS=0; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r03j_bench_2gpu.json 2> gpurun_out/r03j_bench_2gpu.err; echo "bench rc=$?"
echo "wall seconds: $SECONDS"
python tools/show_bench.py gpurun_out/r03j_bench_2gpu.json 2>/dev/null | head -8; tail -5 gpurun_out/r03j_bench_2gpu.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r03j_bench_2gpu.json").readline())
for k, v in d.get("configs", {}).items():
    print(k, {a: (round(b, 1) if isinstance(b, float) else b) for a, b in v.items()})
print("cpu_baseline", d.get("cpu_baseline"))
PY

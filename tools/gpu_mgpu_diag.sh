#!/bin/bash
# multi-GPU schedule diagnostics: usage gpu_mgpu_diag.sh N  (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
run() {
  echo "=== $*"
  env BENCH_VERBOSE=1 "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline 2> gpurun_out/diag.err | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value %.1f (%.1f/GPU) e2e %.1f ms/step %.2f' % (d['value'], d['value']/d['n_gpus'], d['e2e']['value'], d['ms_per_step']))"
  grep "^rank" gpurun_out/diag.err
}
run A=1
run BENCH_PLAIN_SCHEDULE=1
run BENCH_SKIP_NCCL=1

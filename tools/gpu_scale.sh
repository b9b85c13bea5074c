#!/bin/bash
# scaling run on one box: usage gpu_scale.sh N1 N2 ...   (run under gpurun --gpus max(N))
mkdir -p gpurun_out
for N in "$@"; do
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N$N \
      bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$N.json 2> gpurun_out/scale_$N.err
  fi
  echo "N=$N rc=$?"; python -c "
import json
d=json.loads(open('gpurun_out/scale_$N.json').readline()); print('N=%d value %.1f (%.1f/GPU) e2e %.1f ms/step %.2f' % (d['n_gpus'], d['value'], d['value']/d['n_gpus'], d['e2e']['value'], d['ms_per_step']))"
done

#!/bin/bash
# GPU box: quick bench (+ one full-size parity test) of every prebuilt variant under lib/variants/; the default library is restored at the end
PKG=orb-slam-free-space-carving_b200
mkdir -p gpurun_out
cp $PKG/lib/libsdpm.so /tmp/libsdpm.default.so
for d in $PKG/lib/variants/*/; do
  name=$(basename $d)
  cp $d/libsdpm.so $PKG/lib/libsdpm.so
  echo "=== variant $name"
  timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "full_size_720p_n10 or config1_480p" 2>&1 | tail -1
  timeout 600 python bench.py --steps ${VSTEPS:-3} --warmup 3 --no-cpu-baseline --no-configs 2>/dev/null | tee gpurun_out/variant_$name.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value %.1f e2e %.1f  e2e_points %.1f  host_issue %s  standalone_us %s' % (d['value'], d['e2e']['value'], d['e2e_points']['value'], {k: round(v,1) for k,v in d['host_issue_ms_per_step'].items()}, {k: round(v,1) for k,v in d['kernels_standalone_us'].items()}))"
done
cp /tmp/libsdpm.default.so $PKG/lib/libsdpm.so

#!/bin/bash
# bench value for several SDPM_K1_BLOCKS_PER_SM (room left for the post-search kernels that run concurrently)
for b in "$@"; do
  echo "=== SDPM_K1_BLOCKS_PER_SM=$b"
  SDPM_K1_BLOCKS_PER_SM=$b timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('value %.1f e2e %.1f  kernels_avg_us %s' % (d['value'], d['e2e']['value'], {k: round(v,1) for k,v in d['kernels_avg_us'].items()}))"
done

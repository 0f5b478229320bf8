#!/bin/bash
# scratch: one-CTA sort variants on cfg1 (16384 pairs)
mkdir -p gpurun_out
for v in 0 1 2 3 4; do
  FFEMB_BSORT=$v timeout 300 python bench.py --workload 1 --steps 200 --warmup 20 --no-cpu --no-e2e > gpurun_out/v_$v.json 2> gpurun_out/v_$v.err
  python -c "
import json
j=json.loads(open('gpurun_out/v_$v.json').read().strip().splitlines()[-1])
print($v, j['ms_per_step'], j['roofline']['phases_ms_per_step'])
" || tail -3 gpurun_out/v_$v.err
done

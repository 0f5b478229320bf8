#!/bin/bash
# scratch experiment: batch slices of the backward (output-grad slab per table kept inside L2)
mkdir -p gpurun_out
run() { # tag workload slice_mb
  FFEMB_BWD_SLICE_MB=$3 timeout 400 python bench.py --workload $2 --steps 20 --warmup 5 --no-cpu --no-e2e --no-overlap --graph off > gpurun_out/s_$1.json 2> gpurun_out/s_$1.err
  python -c "
import json
j=json.loads(open('gpurun_out/s_$1.json').read().strip().splitlines()[-1])
print('$1', j['ms_per_step'], j['gpu_launches'], j['roofline']['phases_ms_per_step'])
" || tail -3 gpurun_out/s_$1.err
}
run c5_off 5 0
run c5_64 5 64
run c5_32 5 32
run c5_16 5 16
run c3_off 3 0
run c3_16 3 16

#!/bin/bash
# scratch: single-CTA tail threshold scaled with the slots per warp — parity suite, then cfg1 / cfg2 / cfg3
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/pytest_gpu.log
for w in 1 2 3; do
  st=200; [ $w = 3 ] && st=50
  timeout 300 python bench.py --workload $w --steps $st --warmup 10 --no-cpu --no-e2e > gpurun_out/t_cfg${w}.json 2> gpurun_out/t_cfg${w}.err
  python -c "
import json
j=json.loads(open('gpurun_out/t_cfg${w}.json').read().strip().splitlines()[-1])
print($w, j['ms_per_step'], j['value'], j['gpu_launches'], j['roofline']['phases_ms_per_step'])
" || tail -3 gpurun_out/t_cfg${w}.err
done

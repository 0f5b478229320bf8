#!/bin/bash
# scratch experiment: CUDA-graph replay inside HostStep (e2e of the small configs)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "host_step or cuda_graph" > gpurun_out/g_test.log 2>&1
echo "test rc=$?" >> gpurun_out/g_test.log
tail -5 gpurun_out/g_test.log
for w in 1 2; do
  timeout 300 python bench.py --workload $w --steps 200 --warmup 20 --no-cpu > gpurun_out/h_cfg${w}.json 2> gpurun_out/h_cfg${w}.err
done
for f in gpurun_out/h_cfg*.json; do echo $f; python -c "
import json,sys
j=json.loads(open('$f').read().strip().splitlines()[-1])
print(j['ms_per_step'], j['value'], j['gpu_launches'], j['e2e']['ms_per_step'], j['e2e']['value'], j['e2e']['h2d_gbs'], j['e2e']['cuda_graph'])
" || tail -3 ${f%.json}.err; done

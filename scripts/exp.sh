#!/bin/bash
# the other single-GPU BASELINE configs through the full bench line (value, e2e, cpu_baseline)
mkdir -p gpurun_out
for w in 1 2 5; do
  timeout 400 python bench.py --workload $w > gpurun_out/final_cfg$w.json 2> gpurun_out/final_cfg$w.err; echo "cfg$w rc=$?"
  python -c "
import json
d=json.loads(open('gpurun_out/final_cfg$w.json').read().strip().splitlines()[-1])
r=d['roofline']; e=d.get('e2e',{}); c=d.get('cpu_baseline',{})
print('cfg$w', round(d['ms_per_step'],4),'ms', '%.4g'%d['value'], 'graph', d['config']['cuda_graph'], 'frac', r['step_frac_of_peak'], 'e2e_ms', round(e.get('ms_per_step',0),4), '%.4g'%e.get('value',0), 'cpu', '%.4g'%c.get('value',0), c.get('cores'), c.get('kind'))
" || tail -3 gpurun_out/final_cfg$w.err
done

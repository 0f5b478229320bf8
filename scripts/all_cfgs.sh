# every single-GPU BASELINE config through bench.py (device-resident arm only), phases per step
mkdir -p gpurun_out
for w in 1 2 5; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-e2e --no-cpu > gpurun_out/cfg$w.json 2> gpurun_out/cfg$w.err; echo "cfg$w rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/cfg$w.json').read().strip().splitlines()[-1])
r=d['roofline']
print('cfg$w', round(d['ms_per_step'],4),'ms', '%.3g'%d['value'],'lookups/s', 'launches/step', d['gpu_launches']/d['steps'], 'step_frac_peak', r['step_frac_of_peak'], r['phases_ms_per_step'])
"
done

TAG=${1:-r2o}
mkdir -p gpurun_out
for mb in -1 0 32 64 96; do
  python bench.py --steps 20 --warmup 5 --no-cpu --no-e2e --l2-persist $mb > gpurun_out/${TAG}_l2_$mb.json 2> gpurun_out/${TAG}_l2_$mb.err
  grep "l2-persist" gpurun_out/${TAG}_l2_$mb.err
  python -c "
import json
d=json.loads(open('gpurun_out/${TAG}_l2_$mb.json').read().strip().splitlines()[-1])
print('l2-persist $mb: step ms', round(d['ms_per_step'],4), d['roofline']['phases_ms_per_step'])"
done
python bench.py --workload 5 --steps 10 --warmup 3 --no-cpu --no-e2e --l2-persist 64 > gpurun_out/${TAG}_cfg5_l2_64.json 2>/dev/null; python -c "
import json
d=json.loads(open('gpurun_out/${TAG}_cfg5_l2_64.json').read().strip().splitlines()[-1])
print('cfg5 l2-persist 64: step ms', round(d['ms_per_step'],4), d['roofline']['phases_ms_per_step'])"

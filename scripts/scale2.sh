# scaling run (fused exchange, balanced placement) at N = 1,2,4,8 + config 4 row-wise at N = 8
mkdir -p gpurun_out
python bench.py --steps 30 --warmup 5 --no-cpu --no-e2e > gpurun_out/s2_n1.json 2> gpurun_out/s2_n1.err
for N in 2 4 8; do
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
  $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/s2_n$N.json 2> gpurun_out/s2_n$N.err
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29530"
$TR bench.py --gpus 8 --steps 30 --warmup 5 --no-cpu --no-e2e --no-overlap > gpurun_out/s2_n8_noov.json 2> gpurun_out/s2_n8_noov.err
$TR bench.py --gpus 8 --workload 4 --steps 20 --warmup 5 > gpurun_out/s2_cfg4_n8.json 2> gpurun_out/s2_cfg4_n8.err
for f in n1 n2 n4 n8 n8_noov cfg4_n8; do python -c "
import json
try:
    d=json.loads(open('gpurun_out/s2_$f.json').read().strip().splitlines()[-1])
    print('$f', 'ms', round(d['ms_per_step'],3), 'Glookups/s', round(d['value']/1e9,2), d['roofline'].get('phases_ms_per_step'), 'e2e', d.get('e2e',{}).get('ms_per_step'))
except Exception as e: print('$f failed', e)
" 2>/dev/null; done; tail -5 gpurun_out/s2_cfg4_n8.err

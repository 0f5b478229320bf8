# scaling run on one box: bench at N = 1, 2, 4, 8 (or up to $1), fused exchange, both placements at max N
MAXN=${1:-8}
mkdir -p gpurun_out
python bench.py --steps 30 --warmup 5 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for N in 2 4 8; do
  [ $N -le $MAXN ] || continue
  TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N"
  $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
done
N=$MAXN
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29520"
$TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e --placement modulo > gpurun_out/scale_n${N}_modulo.json 2> gpurun_out/scale_n${N}_modulo.err
$TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e --exchange nccl > gpurun_out/scale_n${N}_nccl.json 2> gpurun_out/scale_n${N}_nccl.err
python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -2
for f in n1 n2 n4 n8 n${N}_modulo n${N}_nccl; do python -c "
import json
try:
    d=json.loads(open('gpurun_out/scale_$f.json').read().strip().splitlines()[-1])
    print('$f', 'ms', round(d['ms_per_step'],3), 'Glookups/s', round(d['value']/1e9,2), d['roofline']['phases_ms_per_step'], 'e2e', d.get('e2e',{}).get('ms_per_step'), d['config'].get('table_owner'))
except Exception as e: print('$f failed', e)
" 2>/dev/null; done

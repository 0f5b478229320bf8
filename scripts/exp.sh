N=${1:-2}
python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -15
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu > gpurun_out/bench_n${N}_fused.json 2> gpurun_out/bench_n${N}_fused.err; echo rc=$?
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --exchange nccl > gpurun_out/bench_n${N}_nccl.json 2> gpurun_out/bench_n${N}_nccl.err; echo rc=$?
$TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --placement modulo > gpurun_out/bench_n${N}_modulo.json 2> gpurun_out/bench_n${N}_modulo.err; echo rc=$?
for f in fused nccl modulo; do python -c "
import json,sys
try:
    d=json.loads(open('gpurun_out/bench_n${N}_$f.json').read().strip().splitlines()[-1])
    print('$f', round(d['ms_per_step'],3), d['value']/1e9, d['roofline']['phases_ms_per_step'], 'e2e', d.get('e2e',{}).get('ms_per_step'))
except Exception as e: print('$f failed', e)
"; tail -3 gpurun_out/bench_n${N}_$f.err; done

N=8; TAG=${1:-r2m}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29591"
run() { name=$1; shift
  timeout 400 env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu $ARGS > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_n${N}_$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'parity', d.get('parity_ok'), 'graph', d['config'].get('cuda_graph'), d['roofline'].get('phases_ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/${TAG}_n${N}_$name.err').read()[-2000:])
PY
}
ARGS="" run default A=1
ARGS="--no-e2e" run xpush4 FFEMB_XPUSH=4
ARGS="--workload 4" run cfg4 A=1

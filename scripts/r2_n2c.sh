N=2; TAG=${1:-r2k}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571"
timeout 600 $TR tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n$N.log 2>&1; echo "multigpu_check rc=$?"; grep -i "assert\|ok:\|Error" gpurun_out/${TAG}_mgcheck_n$N.log | head -8
run() { name=$1; shift
  timeout 600 env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu $ARGS > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err; echo "$name rc=$?"
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
ARGS="--no-e2e" run xpush2 FFEMB_XPUSH=2
ARGS="--no-e2e" run xpush8 FFEMB_XPUSH=8
ARGS="--no-e2e" run xpush16 FFEMB_XPUSH=16
ARGS="--no-e2e --workload 4 --rows-cap 20000000" run cfg4_routed A=1

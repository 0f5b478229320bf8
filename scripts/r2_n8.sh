# N = 8 (one box): cfg3 graph / eager / groups / priorities, cfg4 routed vs legacy.  usage: r2_n8.sh TAG
N=8; TAG=${1:-r2j}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561"
run() { name=$1; shift
  timeout 400 env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e $ARGS > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_n${N}_$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'parity', d.get('parity_ok'), 'graph', d['config'].get('cuda_graph'), d['roofline'].get('phases_ms_per_step'))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/${TAG}_n${N}_$name.err').read()[-2000:])
PY
}
ARGS="" run graph_g2 A=1
ARGS="--groups 1" run graph_g1 A=1
ARGS="--groups 1 --no-split" run graph_g1_nosplit A=1
ARGS="" run graph_g2_sideprio0 FFEMB_SIDE_PRIO=0
ARGS="--workload 4" run cfg4_routed A=1
ARGS="--workload 4 --graph off" run cfg4_routed_eager A=1

# round 2, N = $1: multigpu_check (routed row-wise, graph replay), single-GPU emulation test, bench graph on/off, cfg4 (rows capped) routed vs legacy
N=${1:-2}; TAG=${2:-r2e}; CAP=${3:-20000000}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N"
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "rowwise_route" > gpurun_out/${TAG}_emul.log 2>&1; echo "emulated-rank test rc=$?"; tail -3 gpurun_out/${TAG}_emul.log
timeout 600 $TR tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n$N.log 2>&1; echo "multigpu_check rc=$?"; grep -i "assert\|ok:\|Error" gpurun_out/${TAG}_mgcheck_n$N.log | head -8
run() { name=$1; shift
  timeout 600 env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu --no-e2e $ARGS > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_n${N}_$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'parity', d.get('parity_ok'), 'graph', d['config'].get('cuda_graph'), d['roofline'].get('phases_ms_per_step'), (d.get('parity') or {}).get('worst_rel_err'))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/${TAG}_n${N}_$name.err').read()[-2500:])
PY
}
ARGS="" run graph A=1
ARGS="--graph off" run eager A=1
ARGS="--groups 1" run graph_g1 A=1
ARGS="--workload 4 --rows-cap $CAP" run cfg4_routed A=1
ARGS="--workload 4 --rows-cap $CAP --graph off" run cfg4_routed_eager A=1
ARGS="--workload 4 --rows-cap $CAP --rowwise legacy" run cfg4_legacy A=1

# multi-GPU point: parity check (world N) + bench variants at N = $1 GPUs.  Run under gpurun --gpus N.
N=${1:-2}; TAG=${2:-r2b}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N"
timeout 600 $TR tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n$N.log 2>&1; echo "multigpu_check rc=$?"; tail -4 gpurun_out/${TAG}_mgcheck_n$N.log
run() { # name, extra env/args
  name=$1; shift
  timeout 600 env "$@" $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu $ARGS > gpurun_out/${TAG}_n${N}_$name.json 2> gpurun_out/${TAG}_n${N}_$name.err; echo "$name rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_n${N}_$name.json').read().strip().splitlines()[-1])
    print('$name', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,2), 'parity', d.get('parity_ok'), d['roofline'].get('phases_ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'))
except Exception as e:
    print('$name failed', e); print(open('gpurun_out/${TAG}_n${N}_$name.err').read()[-2500:])
PY
}
ARGS="" run default A=1
ARGS="--no-e2e --groups 1" run groups1 A=1
ARGS="--no-e2e" run ldst FFEMB_XPUSH=0
ARGS="--no-e2e --groups 1 --no-split" run r1form FFEMB_XPUSH=0
ARGS="--no-e2e --placement modulo --no-split" run modulo A=1

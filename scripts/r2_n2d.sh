N=2; TAG=${1:-r2l}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29581"
timeout 600 $TR tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n$N.log 2>&1; echo "multigpu_check rc=$?"; grep -i "assert\|ok:\|Error" gpurun_out/${TAG}_mgcheck_n$N.log | head -8
timeout 600 $TR bench.py --gpus $N --steps 30 --warmup 5 --no-cpu > gpurun_out/${TAG}_n${N}_default.json 2> gpurun_out/${TAG}_n${N}_default.err; echo "default rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/${TAG}_n${N}_default.json').read().strip().splitlines()[-1])
print('default', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'parity', d.get('parity_ok'), 'graph', d['config'].get('cuda_graph'), d['roofline'].get('phases_ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'))
PY

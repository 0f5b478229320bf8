TAG=${1:-r2x}
mkdir -p gpurun_out
TR8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29602"
timeout 300 $TR8 tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n8.log 2>&1; echo "multigpu_check rc=$?"; grep -i "assert\|ok:\|Error" gpurun_out/${TAG}_mgcheck_n8.log | head -5
timeout 300 $TR8 bench.py --gpus 8 --steps 30 --warmup 5 --no-cpu --no-e2e > gpurun_out/${TAG}_n8.json 2> gpurun_out/${TAG}_n8.err; echo "n8 rc=$?"
timeout 300 $TR4 bench.py --gpus 4 --steps 30 --warmup 5 --no-cpu --no-e2e > gpurun_out/${TAG}_n4.json 2> gpurun_out/${TAG}_n4.err; echo "n4 rc=$?"
for f in n8 n4; do python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/${TAG}_$f.json').read().strip().splitlines()[-1])
    print('$f', 'ms', round(d['ms_per_step'],4), 'G/s', round(d['value']/1e9,3), 'parity', d.get('parity_ok'), 'graph', d['config'].get('cuda_graph'), d['roofline'].get('phases_ms_per_step'))
except Exception as e:
    print('$f failed', e); print(open('gpurun_out/${TAG}_$f.err').read()[-2000:])
PY
done

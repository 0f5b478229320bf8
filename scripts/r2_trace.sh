N=${1:-2}; TAG=${2:-r2c}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N"
timeout 600 $TR tests/multigpu_check.py > gpurun_out/${TAG}_mgcheck_n$N.log 2>&1; echo "multigpu_check rc=$?"; grep -v "^\s*$" gpurun_out/${TAG}_mgcheck_n$N.log | grep -i "assert\|ok:" | head -5
for G in 2 1; do
timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --no-cpu --no-e2e --trace --groups $G > gpurun_out/${TAG}_n${N}_g$G.json 2> gpurun_out/${TAG}_n${N}_g$G.err; echo "rc=$?"
grep "trace rank" gpurun_out/${TAG}_n${N}_g$G.err
python -c "
import json
d=json.loads(open('gpurun_out/${TAG}_n${N}_g$G.json').read().strip().splitlines()[-1])
print('groups $G ms', round(d['ms_per_step'],4), 'parity', d.get('parity_ok'), d['roofline'].get('phases_ms_per_step'))"
done
